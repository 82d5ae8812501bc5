// oracle.cc — TEST INFRASTRUCTURE ONLY (see oracle.h for scope, status and the
// canonical linearisation rules).  CPU restatement of the reference's
// Filter/Score path; every function cites the reference file:line it follows.
// Build with -ffp-contract=off: Go/amd64 rounds every float32 op separately.
#include "oracle.h"

#include <algorithm>
#include <cassert>
#include <cstdio>
#include <cmath>
#include <cstring>
#include <functional>
#include <set>
#include <stdexcept>
#include <thread>

namespace oracle {

// ---------------------------------------------------------------------------
// Go-semantics helpers
// ---------------------------------------------------------------------------
static inline int32_t i32add(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t i32sub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static inline int32_t i32mul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }

static bool isSpace(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }
static std::string TrimSpace(const std::string& s) {  // strings.TrimSpace (ASCII subset)
  size_t a = 0, b = s.size();
  while (a < b && isSpace((unsigned char)s[a])) a++;
  while (b > a && isSpace((unsigned char)s[b - 1])) b--;
  return s.substr(a, b - a);
}
static std::vector<std::string> Split(const std::string& s, char sep) {  // strings.Split
  std::vector<std::string> out;
  size_t start = 0;
  for (;;) {
    size_t p = s.find(sep, start);
    if (p == std::string::npos) { out.push_back(s.substr(start)); break; }
    out.push_back(s.substr(start, p - start));
    start = p + 1;
  }
  return out;
}
static std::string ToUpper(std::string s) { for (auto& c : s) if (c >= 'a' && c <= 'z') c = (char)(c - 32); return s; }
static std::string ToLower(std::string s) { for (auto& c : s) if (c >= 'A' && c <= 'Z') c = (char)(c + 32); return s; }
static bool Contains(const std::string& s, const std::string& sub) { return s.find(sub) != std::string::npos; }
static bool HasPrefix(const std::string& s, const std::string& p) { return s.compare(0, p.size(), p) == 0; }
static bool ParseBool(const std::string& s, bool& v) {  // strconv.ParseBool
  static const char* T[] = {"1", "t", "T", "TRUE", "true", "True"};
  static const char* F[] = {"0", "f", "F", "FALSE", "false", "False"};
  for (auto t : T) if (s == t) { v = true; return true; }
  for (auto f : F) if (s == f) { v = false; return true; }
  return false;
}
static const std::string* anno(const Pod& pod, const std::string& key) {
  auto it = pod.Annotations.find(key);
  return it == pod.Annotations.end() ? nullptr : &it->second;
}

// ---------------------------------------------------------------------------
// pkg/device/common/common.go:46-53
// ---------------------------------------------------------------------------
std::string GenReason(const std::map<std::string, int>& reasons, int cards) {
  std::vector<std::string> reason;
  for (auto& kv : reasons) {
    char buf[160];
    snprintf(buf, sizeof buf, "%d/%d %s", kv.second, cards, kv.first.c_str());
    reason.push_back(buf);
  }
  std::sort(reason.begin(), reason.end());
  std::string out;
  for (size_t i = 0; i < reason.size(); i++) { if (i) out += ", "; out += reason[i]; }
  return out;
}

// ---------------------------------------------------------------------------
// pkg/device/devices.go:693-738
// ---------------------------------------------------------------------------
bool CheckUUID(const std::map<std::string, std::string>& annos, const std::string& id,
               const std::string& useKey, const std::string& noUseKey) {
  auto match = [&](const std::string& list) {
    for (auto& u : Split(list, ',')) if (TrimSpace(u) == id) return true;
    return false;
  };
  auto it = annos.find(useKey);
  if (it != annos.end() && TrimSpace(it->second) != "") {
    if (!match(it->second)) return false;
  }
  it = annos.find(noUseKey);
  if (it != annos.end() && TrimSpace(it->second) != "") {
    if (match(it->second)) return false;
  }
  return true;
}

bool CheckType(const std::map<std::string, std::string>& annos, const std::string& cardTypeIn,
               const std::string& useKey, const std::string& noUseKey) {
  std::string cardType = ToUpper(cardTypeIn);
  auto match = [&](const std::string& list) {
    for (auto& t0 : Split(list, ',')) {
      std::string t = TrimSpace(t0);
      if (t != "" && Contains(cardType, ToUpper(t))) return true;
    }
    return false;
  };
  auto it = annos.find(useKey);
  if (it != annos.end() && TrimSpace(it->second) != "") {
    if (!match(it->second)) return false;
  }
  it = annos.find(noUseKey);
  if (it != annos.end() && TrimSpace(it->second) != "") {
    if (match(it->second)) return false;
  }
  return true;
}

// pkg/util/util.go:277-298
std::string GetGPUSchedulerPolicyByPod(const std::string& def, const Pod& pod) {
  if (auto v = anno(pod, "hami.io/gpu-scheduler-policy")) return *v;
  return def;
}
bool PolicyContains(const std::string& policy, const std::string& name) {
  for (auto& p : Split(policy, ',')) if (TrimSpace(p) == name) return true;
  return false;
}

// pkg/util/device_scoring_weights.go:58-102
// fmt's %q for the strings an annotation can hold: Go's strconv.Quote (printable text kept, the usual escapes, other bytes \xNN)
static std::string goQuote(const std::string& s) {
  std::string out = "\"";
  for (unsigned char c : s) {
    switch (c) {
      case '"': out += "\\\""; break;
      case '\\': out += "\\\\"; break;
      case '\n': out += "\\n"; break;
      case '\t': out += "\\t"; break;
      case '\r': out += "\\r"; break;
      default:
        if (c < 0x20 || c == 0x7f) { char b[8]; snprintf(b, sizeof b, "\\x%02x", c); out += b; }
        else out += (char)c;
    }
  }
  return out + "\"";
}
static bool ParseDeviceScoringWeightsDetail(const std::string& value, DeviceScoringWeights& weights, std::string& err);
bool ParseDeviceScoringWeights(const std::string& value, DeviceScoringWeights& weights, std::string& err) {
  std::string detail;
  if (ParseDeviceScoringWeightsDetail(value, weights, detail)) return true;
  err = "invalid hami.io/device-scoring-weights annotation " + goQuote(value) + ": " + detail;   // device_scoring_weights.go:62-98
  return false;
}
static bool ParseDeviceScoringWeightsDetail(const std::string& value, DeviceScoringWeights& weights, std::string& err) {
  weights = DeviceScoringWeights{0, 0, 0};
  std::set<std::string> seen;
  auto parts = Split(value, ',');
  if (parts.size() != 3) { err = "expected slot, core, and memory weights"; return false; }
  for (auto& part : parts) {
    std::string p = TrimSpace(part);
    size_t eq = p.find('=');
    if (eq == std::string::npos) { err = "expected key=value entries"; return false; }
    std::string key = TrimSpace(p.substr(0, eq));
    std::string val = TrimSpace(p.substr(eq + 1));
    if (seen.count(key)) { err = "duplicate " + goQuote(key) + " weight"; return false; }
    // strconv.ParseInt(val, 10, 64)
    const std::string notInt = goQuote(key) + " weight must be an integer";
    if (val.empty()) { err = notInt; return false; }
    size_t i = 0; bool neg = false;
    if (val[0] == '+' || val[0] == '-') { neg = val[0] == '-'; i = 1; }
    if (i >= val.size()) { err = notInt; return false; }
    unsigned long long acc = 0;
    for (; i < val.size(); i++) {
      if (val[i] < '0' || val[i] > '9') { err = notInt; return false; }
      unsigned long long d = (unsigned long long)(val[i] - '0');
      if (acc > (0x8000000000000000ULL - d) / 10) { err = notInt; return false; }
      acc = acc * 10 + d;
    }
    if (!neg && acc > 0x7fffffffffffffffULL) { err = notInt; return false; }
    int64_t w = neg ? (int64_t)(0 - acc) : (int64_t)acc;
    if (w < 0) { err = goQuote(key) + " weight must not be negative"; return false; }
    if (key == "slot") weights.Slot = w;
    else if (key == "core") weights.Core = w;
    else if (key == "memory") weights.Memory = w;
    else { err = "unknown weight " + goQuote(key); return false; }
    seen.insert(key);
  }
  if (weights.Slot == 0 && weights.Core == 0 && weights.Memory == 0) { err = "at least one weight must be positive"; return false; }
  return true;
}

// ---------------------------------------------------------------------------
// pkg/scheduler/policy/gpu_policy.go
// ---------------------------------------------------------------------------
// gpu_policy.go:194-222
void DeviceListsScore::ComputeScore(const ContainerDeviceRequests& requests, const DeviceScoringWeights& weights) {
  if (Device.Count == 0 || Device.Totalcore == 0 || Device.Totalmem == 0) { Score = 0; return; }
  int32_t request = 0, core = 0, mem = 0;
  for (auto& kv : requests) {
    const ContainerDeviceRequest& container = kv.second;
    if (container.Type != Device.Type) continue;
    request = i32add(request, 1);
    core = i32add(core, container.Coresreq);
    if (container.MemPercentagereq != 0 && container.MemPercentagereq != 101) {
      mem = i32add(mem, (int32_t)(((int64_t)Device.Totalmem * (int64_t)container.MemPercentagereq) / 100));
      continue;
    }
    mem = i32add(mem, container.Memreq);
  }
  float usedScore = (float)i32add(request, Device.Used) / (float)Device.Count;
  float coreScore = (float)i32add(core, Device.Usedcores) / (float)Device.Totalcore;
  float memScore = (float)i32add(mem, Device.Usedmem) / (float)Device.Totalmem;
  float a = (float)weights.Slot * usedScore;
  float b = (float)weights.Core * coreScore;
  float c = (float)weights.Memory * memScore;
  float s = a + b;
  s = s + c;
  Score = (float)10 * s;
}

// gpu_policy.go:62-75
static std::vector<std::string> gpuSortKeyChain(const std::string& policy) {
  static const char* order[] = {"binpack", "spread", "numa"};
  std::set<std::string> seen;
  std::vector<std::string> chain;
  for (auto& p : Split(policy, ',')) {
    std::string name = TrimSpace(p);
    for (auto key : order) {
      if (name == key && !seen.count(key)) { chain.push_back(key); seen.insert(key); }
    }
  }
  return chain;
}

// gpu_policy.go:77-167
bool DeviceUsageList::Less(int i, int j) const {
  const DeviceListsScore& a = DeviceLists[i];
  const DeviceListsScore& b = DeviceLists[j];
  if (Contains(Policy, ",") || Policy == "numa") {
    // lessByChain, gpu_policy.go:132-167
    std::vector<std::string> chain = gpuSortKeyChain(Policy);
    if (chain.empty()) chain = {"spread"};
    if (NumaBind && chain[0] != "numa") {
      std::vector<std::string> withNuma = {"numa"};
      for (auto& key : chain) if (key != "numa") withNuma.push_back(key);
      chain = withNuma;
    }
    for (auto& key : chain) {
      if (key == "binpack") { if (a.Score != b.Score) return a.Score < b.Score; }
      else if (key == "spread") { if (a.Score != b.Score) return a.Score > b.Score; }
      else if (key == "numa") { if (a.Device.Numa != b.Device.Numa) return a.Device.Numa < b.Device.Numa; }
    }
    return a.Device.Index < b.Device.Index;
  }
  float si = a.Score, sj = b.Score;
  int ni = a.Device.Numa, nj = b.Device.Numa;
  bool binpack = Policy == "binpack";
  if (Policy == "mutex") {
    int32_t ui = a.Device.Used, uj = b.Device.Used;
    if (ui != uj) return ui > uj;
    return ni < nj;
  }
  if (NumaBind) {
    if (binpack) {
      if (ni == nj) return si < sj;
      return ni > nj;
    }
    if (ni == nj) return si > sj;
    return ni < nj;
  }
  if (binpack) {
    if (si != sj) return si < sj;
    return ni < nj;
  }
  if (si != sj) return si > sj;
  return ni < nj;
}

// Go's sort.Sort → pdqsort; for n <= 12 it is insertionSort (stable):
//   for i := a+1; i < b; i++ { for j := i; j > a && Less(j, j-1); j-- { Swap(j, j-1) } }
void DeviceUsageList::Sort() {
  int n = (int)DeviceLists.size();
  if (n > 12) throw std::runtime_error("oracle: >12 devices per node (canonical rule 5)");
  for (int i = 1; i < n; i++)
    for (int j = i; j > 0 && Less(j, j - 1); j--) std::swap(DeviceLists[j], DeviceLists[j - 1]);
}

// ---------------------------------------------------------------------------
// pkg/scheduler/policy/node_policy.go:106-130
// ---------------------------------------------------------------------------
float ComputeDefaultScore(const DeviceUsageList& devices) {
  int32_t used = 0, usedCore = 0, usedMem = 0;
  for (auto& d : devices.DeviceLists) {
    used = i32add(used, d.Device.Used);
    usedCore = i32add(usedCore, d.Device.Usedcores);
    usedMem = i32add(usedMem, d.Device.Usedmem);
  }
  int32_t total = 0, totalCore = 0, totalMem = 0;
  for (auto& d : devices.DeviceLists) {
    total = i32add(total, d.Device.Count);
    totalCore = i32add(totalCore, d.Device.Totalcore);
    totalMem = i32add(totalMem, d.Device.Totalmem);
  }
  if (total == 0 || totalCore == 0 || totalMem == 0) return 0;
  float useScore = (float)used / (float)total;
  float coreScore = (float)usedCore / (float)totalCore;
  float memScore = (float)usedMem / (float)totalMem;
  float s = useScore + coreScore;
  s = s + memScore;
  return (float)10 * s;
}

// ---------------------------------------------------------------------------
// pkg/device/initContainer.go:36-114
// ---------------------------------------------------------------------------
PodDevices CollapseInitContainerUsage(int numInit, const PodDevices& raw) {
  struct usage { int32_t mem = 0, cores = 0, slots = 0; };
  std::map<std::pair<std::string, std::string>, usage> initPeak, appSum;
  for (auto& kv : raw) {
    const std::string& devType = kv.first;
    for (size_t cidx = 0; cidx < kv.second.size(); cidx++) {
      for (auto& dev : kv.second[cidx]) {
        auto key = std::make_pair(devType, dev.UUID);
        if ((int)cidx < numInit) {
          usage& cur = initPeak[key];
          if (dev.Usedmem > cur.mem) cur.mem = dev.Usedmem;
          if (dev.Usedcores > cur.cores) cur.cores = dev.Usedcores;
          cur.slots = 1;
        } else {
          usage& cur = appSum[key];
          cur.mem = i32add(cur.mem, dev.Usedmem);
          cur.cores = i32add(cur.cores, dev.Usedcores);
          cur.slots++;
        }
      }
    }
  }
  PodDevices collapsed;
  for (auto& kv : raw) {
    const std::string& devType = kv.first;
    std::set<std::string> uuidSet;
    for (auto& k : initPeak) if (k.first.first == devType) uuidSet.insert(k.first.second);
    for (auto& k : appSum) if (k.first.first == devType) uuidSet.insert(k.first.second);
    ContainerDevices containerDevs;
    for (auto& uuid : uuidSet) {  // std::set iterates sorted == the reference's sort.Slice by UUID
      usage initU, appU;
      auto a = initPeak.find({devType, uuid});
      if (a != initPeak.end()) initU = a->second;
      auto b = appSum.find({devType, uuid});
      if (b != appSum.end()) appU = b->second;
      ContainerDevice cd;
      cd.UUID = uuid;
      cd.Type = devType;
      cd.Usedmem = std::max(initU.mem, appU.mem);
      cd.Usedcores = std::max(initU.cores, appU.cores);
      cd.Slots = std::max(std::max(initU.slots, appU.slots), 1);
      containerDevs.push_back(cd);
    }
    collapsed[devType] = PodSingleDevice{containerDevs};
  }
  return collapsed;
}

// pkg/device/devices.go:451-484
static std::string EncodeContainerDevices(const ContainerDevices& cd) {
  std::string s;
  for (auto& val : cd) {
    s += val.UUID + "," + val.Type + "," + std::to_string(val.Usedmem) + "," + std::to_string(val.Usedcores) + ":";
  }
  return s;
}
std::string EncodePodSingleDevice(const PodSingleDevice& pd) {
  std::string res;
  for (auto& ctrdevs : pd) { res += EncodeContainerDevices(ctrdevs); res += ";"; }
  return res;
}

// ---------------------------------------------------------------------------
// Vendor Fit — the shared template.  nvidia/device.go:715-876, hygon/device.go:264-369,
// cambricon/device.go:378-488, ascend/device.go:442-617 are one loop with
// per-vendor gates; the gates are spelled out per vendor below.
// ---------------------------------------------------------------------------
bool Devices::AddResourceUsage(DeviceUsage& n, ContainerDevice& ctr) const {
  // nvidia/device.go:645-664 (non-MIG), hygon:257, cambricon:338, ascend:426
  n.Used = i32add(n.Used, 1);
  n.Usedcores = i32add(n.Usedcores, ctr.Usedcores);
  n.Usedmem = i32add(n.Usedmem, ctr.Usedmem);
  return true;
}

// ---- kunlun/topo.go -----------------------------------------------------------------------------
static bool klCanMeet(const std::vector<int>& have, const std::vector<int>& want) {       // :163-175
  for (int v : want) if (std::find(have.begin(), have.end(), v) == have.end()) return false;
  return true;
}
static std::vector<int> klDevicePick(const std::vector<bool>& fit, const std::vector<int>& index, int start, int count) {   // :192-204
  std::vector<int> res;
  for (int t = start; t < 8; t++) {
    if (t >= (int)fit.size()) throw std::runtime_error("kunlun devicepick past the end of the device list (the reference panics)");
    if (fit[t]) { res.push_back(index[t]); if ((int)res.size() == count) return res; }
  }
  return res;
}
static std::vector<int> klInterconnect(const std::vector<bool>& fit, const std::vector<int>& index, int count) {   // :112-161
  static const int inter[4][2] = {{0, 4}, {1, 5}, {2, 6}, {3, 7}};                             // InterGroupConnection
  static const int inter2[6][4] = {{0, 1, 4, 5}, {2, 3, 6, 7}, {0, 2, 4, 6}, {1, 3, 5, 7}, {0, 3, 4, 7}, {1, 2, 5, 6}};   // InterGroupConnection2
  static const int group[12][2] = {{0, 1}, {0, 2}, {0, 3}, {1, 2}, {1, 3}, {2, 3}, {4, 5}, {4, 6}, {4, 7}, {5, 6}, {5, 7}, {6, 7}};   // GroupConnection
  const size_t n = fit.size();
  if (count == 2) {
    for (size_t a = 0; a < n; a++) {
      if (!fit[a]) continue;
      for (size_t b = 0; b < n; b++) {
        if (!fit[b] || index[b] == index[a]) continue;
        for (auto& p : inter)
          if ((p[0] == index[a] && p[1] == index[b]) || (p[0] == index[b] && p[1] == index[a])) return {index[a], index[b]};
      }
    }
  }
  if (count == 4) {
    std::vector<int> unused;                                                                    // parseUsage, :30-38
    for (size_t i = 0; i < n; i++) if (fit[i]) unused.push_back(index[i]);
    if (unused.size() == 4 || unused.size() == 5)
      for (auto& c : inter2) { std::vector<int> cv(c, c + 4); if (klCanMeet(unused, cv)) return cv; }
    if (unused.size() == 6) {
      std::vector<int> ret;
      for (auto& c : inter2) {
        std::vector<int> cv(c, c + 4);
        if (!klCanMeet(unused, cv)) continue;
        ret = cv;
        std::vector<int> d;                                                                     // delta, :177-190
        for (int v : unused) if (std::find(cv.begin(), cv.end(), v) == cv.end()) d.push_back(v);
        for (auto& p : inter) if (klCanMeet(d, {p[0], p[1]})) return ret;                       // parseInterconnection: inter-group pairs, then in-group pairs
        for (auto& p : group) if (klCanMeet(d, {p[0], p[1]})) return ret;
      }
      return ret;
    }
  }
  return {};
}
std::vector<int> KunlunGraphSelect(const std::vector<bool>& fit, const std::vector<int>& index, int count) {   // :206-252
  int leftwing = 0, rightwing = 0;
  for (size_t idx = 0; idx < fit.size(); idx++) {
    if (idx < 4) { if (fit[idx]) leftwing++; }
    else if (fit[idx]) rightwing++;
  }
  static const int oddorder[4] = {1, 3, 2, 4};
  if (count == 8) {
    if (leftwing + rightwing == count) return {0, 1, 2, 3, 4, 5, 6, 7};
    return {};
  }
  if (count == 1 || count == 2 || count == 4) {
    if (leftwing >= count || rightwing >= count)
      for (int slots = count; slots <= 4; slots++) {
        int num = slots;
        if (count % 2 == 1) num = oddorder[slots - 1];
        if (leftwing == num) return klDevicePick(fit, index, 0, count);
        if (rightwing == num) return klDevicePick(fit, index, 4, count);
      }
    return klInterconnect(fit, index, count);
  }
  return {};
}
float KunlunCalcScore(const std::vector<int>& p, const std::vector<int>& c) {                   // :74-95 (the sorts compare positions: no-ops)
  auto getvalue = [](int t) { return t == 4 ? 0 : (t == 1 ? 2 : 1); };                          // :50-58
  auto countbubble = [&](const std::vector<int>& t) {                                           // :60-72
    int left = 0, right = 0;
    for (int v : t) { if (v < 4) left++; else right++; }
    if (left == 0 && right == 0) return 1;
    return getvalue(left) + getvalue(right);
  };
  switch (countbubble(c) - countbubble(p)) {
    case -1: return 3000;
    case 0: return 2000;
    case 1: return 1000;
    case 2: return 0;
  }
  return 1000;
}

// ---- dynamic MIG: nvidia/mig_topology.go ---------------------------------------------------
static bool migPlacementsOverlap(const MigPlacement& a, const MigPlacement& b) {   // mig_topology.go:78-80
  return (uint64_t)a.Start < (uint64_t)b.Start + b.Size && (uint64_t)b.Start < (uint64_t)a.Start + a.Size;
}
static std::vector<MigPlacement> orderedMigPlacements(const std::string& profile, std::vector<MigPlacement> ordered) {   // :30-40
  const bool high = HasPrefix(profile, "1g.") || HasPrefix(profile, "3g.");    // preferHighMigPlacement, :26-28
  std::stable_sort(ordered.begin(), ordered.end(), [&](const MigPlacement& a, const MigPlacement& b) {
    return high ? a.Start > b.Start : a.Start < b.Start;
  });
  return ordered;
}
bool SelectMigPlacement(const std::vector<MigProfile>& profiles, const std::vector<MigPlacement>& occupied,
                        const std::string& profileName, MigPlacement& out) {       // :57-76
  const MigProfile* profile = nullptr;
  for (auto& p : profiles) if (p.Name == profileName) { profile = &p; break; }    // findMigProfile, :42-49
  if (!profile) return false;
  for (auto& candidate : orderedMigPlacements(profileName, profile->Placements)) {
    bool conflict = false;
    for (auto& existing : occupied) if (migPlacementsOverlap(candidate, existing)) { conflict = true; break; }
    if (!conflict) { out = candidate; return true; }
  }
  return false;
}
bool SelectMigCandidate(const std::vector<MigProfile>& profiles, const std::vector<MigPlacement>& occupied, int32_t memory,
                        MigProfile& profile, MigPlacement& placement) {           // nvidia/device.go:621-639
  std::vector<MigProfile> candidates = profiles;
  std::stable_sort(candidates.begin(), candidates.end(), [](const MigProfile& a, const MigProfile& b) {
    if (a.MemoryMB != b.MemoryMB) return a.MemoryMB < b.MemoryMB;
    return a.SliceCount < b.SliceCount;
  });
  for (auto& c : candidates) {
    if (c.MemoryMB < memory) continue;
    if (SelectMigPlacement(profiles, occupied, c.Name, placement)) { profile = c; return true; }
  }
  return false;
}
bool CanPlaceMigProfiles(const std::vector<MigProfile>& capabilities, const std::vector<MigPlacement>& occupied,
                         const std::vector<std::string>& requested) {             // mig_topology.go:85-132
  if (requested.empty()) return true;
  std::map<std::string, std::vector<MigPlacement>> placements;
  for (auto& p : capabilities) placements[p.Name] = p.Placements;
  std::vector<std::string> ordered = requested;
  std::stable_sort(ordered.begin(), ordered.end(), [&](const std::string& a, const std::string& b) {
    const auto& left = placements[a]; const auto& right = placements[b];
    if (left.size() != right.size()) return left.size() < right.size();
    if (left.empty()) return false;
    return left[0].Size > right[0].Size;
  });
  std::vector<MigPlacement> used = occupied;
  std::function<bool(size_t)> place = [&](size_t idx) {
    if (idx == ordered.size()) return true;
    for (auto& candidate : orderedMigPlacements(ordered[idx], placements[ordered[idx]])) {
      bool conflict = false;
      for (auto& existing : used) if (migPlacementsOverlap(candidate, existing)) { conflict = true; break; }
      if (conflict) continue;
      used.push_back(candidate);
      if (place(idx + 1)) return true;
      used.pop_back();
    }
    return false;
  };
  return place(0);
}
static std::vector<MigPlacement> occupiedMigPlacements(const std::vector<MigAllocation>& allocations) {   // :51-57
  std::vector<MigPlacement> out;
  for (auto& a : allocations) out.push_back(a.Placement);
  return out;
}
// EncodeMigAllocations, nvidia/mig_allocations.go:48-76 (encoding/json: struct field order, omitempty runtime fields absent)
static std::string jsonString(const std::string& v) {
  std::string o = "\"";
  for (unsigned char c : v) {
    if (c == '"' || c == '\\') { o += '\\'; o += (char)c; }
    else if (c == '<' || c == '>' || c == '&') { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }   // encoding/json escapes HTML
    else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
    else o += (char)c;
  }
  return o + "\"";
}
std::string EncodeMigAllocations(const PodSingleDevice& pd) {
  std::string out;
  int n = 0;
  for (size_t containerIndex = 0; containerIndex < pd.size(); containerIndex++)
    for (size_t deviceIndex = 0; deviceIndex < pd[containerIndex].size(); deviceIndex++) {
      const ContainerDevice& dev = pd[containerIndex][deviceIndex];
      if (!dev.HasMig || dev.MigProfileName.empty() || dev.MigPlace.Size == 0) continue;
      out += n++ ? "," : "";
      out += "{\"containerIndex\":" + std::to_string(containerIndex) + ",\"deviceIndex\":" + std::to_string(deviceIndex) +
             ",\"gpuUUID\":" + jsonString(dev.UUID) + ",\"profile\":" + jsonString(dev.MigProfileName) +
             ",\"placement\":{\"start\":" + std::to_string(dev.MigPlace.Start) + ",\"size\":" + std::to_string(dev.MigPlace.Size) + "}}";
    }
  return n ? "[" + out + "]" : "";
}

namespace {

// PLAIN: iluvatar's Fit (iluvatar/device.go:267-371) = the hygon gate list with checkType "key equals the word";
// COUNT_ONLY: biren's Fit (biren/device.go:190-248) = the gates up to mutex, Usedmem 0, Usedcores as requested
// AMD: amd/device.go:279-371 — type lists like hygon, whole-card memory by default, cores as a share of the card
// MTHREADS: mthreads/device.go:289-405 — the template gates plus customFilterRule (all containers of a pod share one card)
// WHOLE_CARD: enflame GCU (enflame/gcu.go:144-183) — health, type, "in use at all" = conflict; no uuid lists, no slot count
// METAX: the whole-GPU "Metax-GPU" device (metax/device.go:223-334) — the template gates with no uuid lists
//   (checkUUID is constant true, :124-126) plus the same customFilterRule as mthreads (:156-166)
enum class Vendor { NVIDIA, DCU, MLU, ASCEND, PLAIN, COUNT_ONLY, AMD, MTHREADS, WHOLE_CARD, METAX, VASTAI };

struct FitGates {
  Vendor vendor;
  std::string commonWord;
  std::string useType, noUseType;   // type selector annotation keys ("" = none)
  std::string useUUID, noUseUUID;
  bool ascendHamiVnpuCoreDefault = false;
};

// generateCombinations, nvidia/device.go:887-915 (lexicographic index order)
void combos(const ContainerDevices& arr, int start, int k, ContainerDevices& cur, std::vector<ContainerDevices>& out) {
  if (k == 0) { out.push_back(cur); return; }
  for (int i = start; i <= (int)arr.size() - k; i++) {
    cur.push_back(arr[i]);
    combos(arr, i + 1, k - 1, cur, out);
    cur.pop_back();
  }
}

int pairScore(const NodeInfo* nodeInfo, const std::string& a, const std::string& b) {
  // getDevicePairScoreMap, nvidia/device.go:917-925; missing entries read as 0 (Go map zero value)
  if (!nodeInfo) return 0;
  auto it = nodeInfo->Devices.find("NVIDIA");
  if (it == nodeInfo->Devices.end()) return 0;
  for (auto& d : it->second) {
    if (d.ID == a) {
      auto s = d.PairScores.find(b);
      return s == d.PairScores.end() ? 0 : s->second;
    }
  }
  return 0;
}

// computeWorstSingleCard, nvidia/device.go:927-951
ContainerDevices computeWorstSingleCard(const NodeInfo* nodeInfo, const ContainerDevices& devices) {
  int worstScore = 0;
  ContainerDevices worst;
  bool found = false;
  for (auto& dev1 : devices) {
    int total = 0;
    for (auto& dev2 : devices) {
      if (dev1.UUID == dev2.UUID) continue;
      total += pairScore(nodeInfo, dev1.UUID, dev2.UUID);
    }
    if (!found || total < worstScore) { found = true; worstScore = total; worst = ContainerDevices{dev1}; }
  }
  return worst;
}

// computeBestCombination, nvidia/device.go:953-978
ContainerDevices computeBestCombination(const NodeInfo* nodeInfo, const std::vector<ContainerDevices>& combinations) {
  int bestScore = -1;
  ContainerDevices best;
  bool found = false;
  for (auto& partition : combinations) {
    int total = 0;
    for (size_t i = 0; i + 1 < partition.size(); i++)
      for (size_t z = i + 1; z < partition.size(); z++)
        total += pairScore(nodeInfo, partition[i].UUID, partition[z].UUID);
    if (!found || total > bestScore) { found = true; bestScore = total; best = partition; }
  }
  return best;
}

// ascend/device.go:631-675 (canonical rule 7 for equal counts)
ContainerDevices ascendBestCombination(int reqNum, const ContainerDevices& containerDevices) {
  std::map<int, ContainerDevices> networkDeviceMap;
  for (auto& cd : containerDevices)
    if (cd.HasNetworkID) networkDeviceMap[cd.NetworkID].push_back(cd);
  std::vector<std::pair<int, int>> sortedNetworks;  // (NetworkID, Count), ascending NetworkID from std::map
  for (auto& kv : networkDeviceMap) sortedNetworks.push_back({kv.first, (int)kv.second.size()});
  std::stable_sort(sortedNetworks.begin(), sortedNetworks.end(),
                   [](const std::pair<int, int>& a, const std::pair<int, int>& b) { return a.second > b.second; });
  ContainerDevices result;
  for (auto& item : sortedNetworks) {
    for (auto& dev : networkDeviceMap[item.first]) {
      result.push_back(dev);
      if ((int)result.size() == reqNum) return result;
    }
  }
  return result;
}

// ascend/device.go:677-733 (canonical rule 7 for equal sizes)
ContainerDevices ascendBestCombination910C(int reqNum, const ContainerDevices& containerDevices) {
  std::map<int, ContainerDevice> indexToDevice;
  std::vector<int> npuIndices;
  for (auto& dev : containerDevices) { indexToDevice[dev.Idx] = dev; npuIndices.push_back(dev.Idx); }
  const int MaxCardNPUNum = 2;
  std::map<int, std::vector<int>> cardTopology;  // ascending card id
  for (int idx : npuIndices) cardTopology[idx / MaxCardNPUNum].push_back(idx);
  std::vector<std::vector<int>> cardTopSlice;
  for (auto& kv : cardTopology) cardTopSlice.push_back(kv.second);
  std::stable_sort(cardTopSlice.begin(), cardTopSlice.end(),
                   [](const std::vector<int>& a, const std::vector<int>& b) { return a.size() > b.size(); });
  std::vector<int> selected;
  int taskNPUNum = reqNum;
  for (auto& card : cardTopSlice) {
    if (taskNPUNum <= 0) break;
    if ((int)card.size() == MaxCardNPUNum) {
      selected.insert(selected.end(), card.begin(), card.end());
      taskNPUNum -= MaxCardNPUNum;
    }
  }
  ContainerDevices result;
  for (int idx : selected) {
    auto it = indexToDevice.find(idx);
    if (it != indexToDevice.end()) result.push_back(it->second);
  }
  return result;
}

// fitQuota, nvidia/device.go:666-692 and cambricon/device.go:351-376: the pod's hypothetical total (containers placed so
// far + this container's picks + the candidate device), collapsed, against the namespace quota
bool fitQuota(const QuotaManager& qm, const ResourceNames& names, const Pod& pod, const ContainerDevices& tmpDevs,
              const PodDevices* allocated, const std::string& word, const std::string& devUUID, int64_t memreq, int64_t coresreq) {
  PodDevices hypo;
  if (allocated) hypo = *allocated;
  ContainerDevices cur = tmpDevs;
  ContainerDevice cd;
  cd.UUID = devUUID; cd.Type = word; cd.Usedmem = (int32_t)memreq; cd.Usedcores = (int32_t)coresreq;
  cur.push_back(cd);
  hypo[word].push_back(cur);
  int64_t mem = 0, core = 0;
  PodDevices collapsed = CollapseInitContainerUsage(pod.NumInitContainers, hypo);
  for (auto& ctr : collapsed[word]) for (auto& v : ctr) { mem += v.Usedmem; core += v.Usedcores; }
  return qm.FitQuota(pod.Namespace, mem, names, core);
}

FitResult fitTemplate(const FitGates& g, const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request,
                      const Pod& pod, const NodeInfo* nodeInfo, const std::string& gpuPolicyDefault,
                      const PodDevices* allocated, const QuotaManager* quota, const ResourceNames& names) {
  ContainerDeviceRequest k = request;
  int32_t originReq = k.Nums;
  int prevnuma = -1;
  ContainerDevices tmpDevs;
  std::map<std::string, int> reason;
  std::string gpuPolicy = GetGPUSchedulerPolicyByPod(gpuPolicyDefault, pod);
  bool isMutex = PolicyContains(gpuPolicy, "mutex");
  bool needTopology = false;
  bool isHAMiCore = false;

  std::set<std::string> cordoned;
  if (g.vendor == Vendor::NVIDIA) {
    needTopology = PolicyContains(gpuPolicy, "topology-aware");  // nvidia/device.go:724
    // cordonedDevices, nvidia/device.go:698-713
    if (nodeInfo && nodeInfo->DeviceCordon && TrimSpace(*nodeInfo->DeviceCordon) != "") {
      for (auto& u0 : Split(*nodeInfo->DeviceCordon, ',')) {
        std::string u = TrimSpace(u0);
        if (u != "") cordoned.insert(u);
      }
    }
  }
  if (g.vendor == Vendor::ASCEND) {
    // ascend/device.go:452-480
    std::string vnpuMode;
    if (auto v = anno(pod, "huawei.com/vnpu-mode")) vnpuMode = *v;
    isHAMiCore = vnpuMode == "hami-core";
    bool nodeSupportHamiCore = g.ascendHamiVnpuCoreDefault;
    if (nodeInfo && nodeInfo->HamiVnpuCore) nodeSupportHamiCore = *nodeInfo->HamiVnpuCore == "true";
    if (isHAMiCore && !nodeSupportHamiCore) {
      reason["ModeNotFit"]++;
      FitResult r; r.fit = false; r.reason = GenReason(reason, (int)devices.size());
      return r;
    }
    if (HasPrefix(g.commonWord, "Ascend910")) {
      bool all = true;  // hasNetworkID, ascend/device.go:619-629
      for (auto d : devices) if (!d->HasNetworkID) { all = false; break; }
      if (all) needTopology = true;
    }
  }
  const bool is910CMulti = g.vendor == Vendor::ASCEND && k.Type == "Ascend910C" && originReq > 1;
  // vastai: isDieMode looks at the FIRST device of the (sorted) list (vastai/device.go:389-403); occupied cards under mutex (:241-250)
  const bool vastaiDie = g.vendor == Vendor::VASTAI && !devices.empty() && devices[0]->HasNetworkID;
  std::set<int> vastaiBusyAIC;
  if (vastaiDie && isMutex)
    for (auto d : devices) if (d->Used > 0 && d->HasNetworkID) vastaiBusyAIC.insert(d->NetworkID);

  for (int i = (int)devices.size() - 1; i >= 0; i--) {
    DeviceUsage* dev = devices[i];
    if (!dev->Health) { reason["CardNotHealth"]++; continue; }
    if (g.vendor == Vendor::NVIDIA && cordoned.count(dev->ID)) { reason["CardCordoned"]++; continue; }
    // checkType: nvidia/device.go:493-503, hygon:161-166, cambricon:250-255, ascend:303-308
    bool found = false, numa = false;
    if (k.Type == g.commonWord) {
      switch (g.vendor) {
        case Vendor::NVIDIA: {
          bool typeCheck = CheckType(pod.Annotations, dev->Type, g.useType, g.noUseType);
          if (auto mode = anno(pod, "nvidia.com/vgpu-mode")) {
            if (!Contains(*mode, dev->Mode)) typeCheck = false;
          }
          found = typeCheck;
          bool enforce = false;  // assertNuma, nvidia/device.go:482-491
          if (auto nb = anno(pod, "nvidia.com/numa-bind")) { bool v; if (ParseBool(*nb, v) && v) enforce = true; }
          numa = enforce;
          break;
        }
        case Vendor::DCU:
        case Vendor::AMD:          // checkAMDType, amd/device.go:165-186: the same two lists
          found = CheckType(pod.Annotations, dev->Type, g.useType, g.noUseType);
          break;
        case Vendor::MLU:
        case Vendor::ASCEND:
        case Vendor::WHOLE_CARD:   // enflame/gcu.go:202-204
        case Vendor::MTHREADS:     // mthreads/device.go:191-196
        case Vendor::METAX:        // metax/device.go:117-122
        case Vendor::PLAIN:        // iluvatar/device.go:190-195
        case Vendor::COUNT_ONLY:   // biren/device.go:130-135
        case Vendor::VASTAI:       // vastai/device.go:126-131
          found = true;
          break;
      }
    }
    if (!found) { reason["CardTypeMismatch"]++; continue; }
    if (g.vendor == Vendor::WHOLE_CARD) {     // enflame/gcu.go:159-176
      if (dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
      if (k.Nums > 0) {
        k.Nums--;
        ContainerDevice cd;
        cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = k.Type; cd.Usedmem = k.Memreq; cd.Usedcores = k.Coresreq;
        tmpDevs.push_back(cd);
      }
      if (k.Nums == 0) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
      continue;
    }
    if (numa && prevnuma != dev->Numa) {
      if (k.Nums != originReq) reason["NumaNotFit"] += (int)tmpDevs.size();
      k.Nums = originReq;
      prevnuma = dev->Numa;
      tmpDevs.clear();
    }
    if (g.vendor != Vendor::METAX && !CheckUUID(pod.Annotations, dev->ID, g.useUUID, g.noUseUUID)) { reason["CardUuidMismatch"]++; continue; }

    int32_t memreq = 0;
    if (dev->Count <= dev->Used) { reason["CardTimeSlicingExhausted"]++; continue; }
    if (isMutex && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
    if (g.vendor == Vendor::VASTAI) {
      // vastai/device.go:228-327.  Card mode is biren's gate list; in DIE mode (the first device of the list says
      // DeviceStrategy "die": CustomInfo["AIC"] rides in NetworkID) a card is made of several dies: under mutex an idle
      // die on a card with a die in use is a conflict too, every fitting die is collected, and the choice goes by card
      if (isMutex && vastaiDie && dev->HasNetworkID && vastaiBusyAIC.count(dev->NetworkID)) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
      if (k.Nums > 0) {
        if (!vastaiDie) k.Nums--;
        ContainerDevice cd;
        cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = k.Type; cd.Usedmem = 0; cd.Usedcores = k.Coresreq;
        cd.HasNetworkID = dev->HasNetworkID; cd.NetworkID = dev->NetworkID;      // CustomInfo: dev.CustomInfo
        tmpDevs.push_back(cd);
      }
      if (k.Nums == 0 && !vastaiDie) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
      continue;
    }
    if (g.vendor == Vendor::COUNT_ONLY) {   // biren/device.go:226-241: no memory / core gates, the pick carries no memory
      if (k.Nums > 0) {
        k.Nums--;
        ContainerDevice cd;
        cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = k.Type; cd.Usedmem = 0; cd.Usedcores = k.Coresreq;
        tmpDevs.push_back(cd);
      }
      if (k.Nums == 0) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
      continue;
    }
    if (g.vendor == Vendor::AMD) {          // amd/device.go:308-350: whole card by default, cores as a share of the card
      int32_t memReq = k.Memreq;
      if (memReq <= 0 && dev->Totalmem > 0) memReq = dev->Totalmem;
      if (i32sub(dev->Totalmem, dev->Usedmem) < memReq) { reason["CardInsufficientMemory"]++; continue; }
      int32_t coreReq = 0;
      if (k.Coresreq > 0) {
        if (dev->Totalcore <= 0) { reason["CardInsufficientCore"]++; continue; }
        coreReq = i32mul(dev->Totalcore, k.Coresreq) / 100;
        coreReq = std::max(coreReq, (int32_t)1);
        coreReq = std::min(coreReq, dev->Totalcore);
      } else if (dev->Totalmem > 0 && memReq >= dev->Totalmem) {
        coreReq = dev->Totalcore;
      }
      if (i32sub(dev->Totalcore, dev->Usedcores) < coreReq) { reason["CardInsufficientCore"]++; continue; }
      if (k.Nums > 0) {
        k.Nums--;
        ContainerDevice cd;
        cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = dev->Type;   // the registered product type, not k.Type (:341-346)
        cd.Usedmem = memReq; cd.Usedcores = coreReq;
        tmpDevs.push_back(cd);
      }
      if (k.Nums == 0) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
      continue;
    }
    if (k.Coresreq > 100) k.Coresreq = 100;
    if (k.Memreq > 0) memreq = k.Memreq;
    if (k.MemPercentagereq != 101 && k.Memreq == 0) memreq = i32mul(dev->Totalmem, k.MemPercentagereq) / 100;
    // nvidia/device.go:784-788, cambricon/device.go:439-443 (hygon and ascend Fit do not consult the quota)
    if ((g.vendor == Vendor::NVIDIA || g.vendor == Vendor::MLU) && quota &&
        !fitQuota(*quota, names, pod, tmpDevs, allocated, g.vendor == Vendor::NVIDIA ? "NVIDIA" : "MLU", dev->ID, memreq, k.Coresreq)) {
      reason["ResourceQuotaNotFit"]++;
      continue;
    }
    if (i32sub(dev->Totalmem, dev->Usedmem) < memreq) { reason["CardInsufficientMemory"]++; continue; }
    int32_t effectiveTotalCore = dev->Totalcore;
    if (isHAMiCore) effectiveTotalCore = 100;  // ascend/device.go:540-543
    if (i32sub(effectiveTotalCore, dev->Usedcores) < k.Coresreq) { reason["CardInsufficientCore"]++; continue; }
    if (effectiveTotalCore == 100 && k.Coresreq == 100 && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
    if (effectiveTotalCore != 0 && dev->Usedcores == effectiveTotalCore && k.Coresreq == 0) { reason["CardComputeUnitsExhausted"]++; continue; }
    if (g.vendor == Vendor::NVIDIA && dev->Mode == "mig") {
      // CustomFilterRule, nvidia/device.go:602-619, with the resolved memory request (:811-814): the picks this Fit
      // already queued on the same GPU take their placements first
      std::vector<MigPlacement> occupied = occupiedMigPlacements(dev->MigAllocationsInUse);
      bool ok = true;
      MigProfile prof; MigPlacement place;
      for (auto& existing : tmpDevs) {
        if (existing.UUID != dev->ID) continue;
        if (!SelectMigCandidate(dev->MigProfiles, occupied, existing.Usedmem, prof, place)) { ok = false; break; }
        occupied.push_back(place);
      }
      if (ok) ok = SelectMigCandidate(dev->MigProfiles, occupied, memreq, prof, place);
      if (!ok) { reason["CardMigTopologyInfeasible"]++; continue; }   // :819-823
    }
    if ((g.vendor == Vendor::MTHREADS || g.vendor == Vendor::METAX) && allocated) {
      // customFilterRule, mthreads/device.go:266-276,362-366: every device an earlier container of this pod got (looked
      // up under the DEVICE's type string) must be this very device
      bool ok = true;
      auto prev = allocated->find(dev->Type);
      if (prev != allocated->end())
        for (auto& ctrs : prev->second) for (auto& cd : ctrs) if (cd.UUID != dev->ID) ok = false;
      if (!ok) { reason["CardNotFoundCustomFilterRule"]++; continue; }
    }

    bool decrement;
    if (g.vendor == Vendor::NVIDIA) decrement = !needTopology;                               // nvidia/device.go:831-833
    else if (g.vendor == Vendor::ASCEND) decrement = !needTopology && !is910CMulti;          // ascend/device.go:564
    else decrement = true;
    if (k.Nums > 0) {
      if (decrement) k.Nums--;
      ContainerDevice cd;
      cd.Idx = (int)dev->Index;
      cd.UUID = dev->ID;
      cd.Type = k.Type;
      cd.Usedmem = memreq;
      cd.Usedcores = k.Coresreq;
      if (g.vendor == Vendor::ASCEND) { cd.HasNetworkID = dev->HasNetworkID; cd.NetworkID = dev->NetworkID; }
      tmpDevs.push_back(cd);
    }
    if (k.Nums == 0 && decrement) {
      FitResult r; r.fit = true; r.devs = tmpDevs;
      return r;
    }
    if (g.vendor == Vendor::NVIDIA && dev->Mode == "mig") i++;   // nvidia/device.go:846-848: the same GPU is visited again
  }

  if (vastaiDie) {    // vastai/device.go:306-320; computeBestCombination :336-372 (cards with equal die counts: ascending AIC, canonical rule 7)
    if ((int)tmpDevs.size() == (int)originReq) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
    if ((int)tmpDevs.size() > (int)originReq) {
      FitResult r; r.fit = true;
      if (originReq == 1) r.devs = ContainerDevices{tmpDevs[0]};
      else r.devs = ascendBestCombination((int)originReq, tmpDevs);
      return r;
    }
  }
  if (is910CMulti) {  // ascend/device.go:582-593
    ContainerDevices combination = ascendBestCombination910C((int)originReq, tmpDevs);
    if ((int)combination.size() != (int)originReq) {
      reason["AllocatedCardsInsufficientRequest"] = (int)combination.size();
      FitResult r; r.fit = false; r.devs = tmpDevs; r.reason = GenReason(reason, (int)originReq);
      return r;
    }
    FitResult r; r.fit = true; r.devs = combination;
    return r;
  }
  if (needTopology) {
    if ((int)tmpDevs.size() == (int)originReq) { FitResult r; r.fit = true; r.devs = tmpDevs; return r; }
    if ((int)tmpDevs.size() > (int)originReq) {
      FitResult r; r.fit = true;
      if (g.vendor == Vendor::NVIDIA) {  // nvidia/device.go:855-868
        if (originReq == 1) r.devs = computeWorstSingleCard(nodeInfo, tmpDevs);
        else {
          std::vector<ContainerDevices> combinations;
          ContainerDevices cur;
          combos(tmpDevs, 0, (int)request.Nums, cur, combinations);
          r.devs = computeBestCombination(nodeInfo, combinations);
        }
      } else {  // ascend/device.go:599-607
        if (originReq == 1) r.devs = ContainerDevices{tmpDevs[0]};
        else r.devs = ascendBestCombination((int)originReq, tmpDevs);
      }
      return r;
    }
  }
  if (!tmpDevs.empty()) reason["AllocatedCardsInsufficientRequest"] = (int)tmpDevs.size();
  FitResult r; r.fit = false; r.devs = tmpDevs; r.reason = GenReason(reason, (int)devices.size());
  return r;
}

class NvidiaGPUDevices : public Devices {
 public:
  std::string CommonWord() const override { return "NVIDIA"; }
  FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                const NodeInfo* nodeInfo, const std::string& def, const PodDevices* allocated,
                const QuotaManager* quota) const override {
    FitGates g{Vendor::NVIDIA, "NVIDIA", "nvidia.com/use-gputype", "nvidia.com/nouse-gputype",
               "nvidia.com/use-gpuuuid", "nvidia.com/nouse-gpuuuid"};
    return fitTemplate(g, devices, request, pod, nodeInfo, def, allocated, quota, GetResourceNames());
  }
  bool AddResourceUsage(DeviceUsage& n, ContainerDevice& ctr) const override {   // nvidia/device.go:645-664
    if (n.Mode == "mig") {
      MigProfile profile; MigPlacement placement;
      if (!SelectMigCandidate(n.MigProfiles, occupiedMigPlacements(n.MigAllocationsInUse), ctr.Usedmem, profile, placement))
        return false;   // "MIG profile and placement allocation failed"
      ctr.Usedmem = profile.MemoryMB;
      ctr.Usedcores = profile.Core;
      ctr.HasMig = true; ctr.MigProfileName = profile.Name; ctr.MigPlace = placement;
      n.MigAllocationsInUse.push_back(MigAllocation{profile.Name, placement});
    }
    return Devices::AddResourceUsage(n, ctr);
  }
  void PatchAnnotations(const Pod&, std::map<std::string, std::string>& annos, const PodDevices& pd) const override {
    // nvidia/device.go:505-518
    auto it = pd.find("NVIDIA");
    if (it != pd.end() && !it->second.empty()) {
      std::string s = EncodePodSingleDevice(it->second);
      annos["hami.io/vgpu-devices-to-allocate"] = s;
      annos["hami.io/vgpu-devices-allocated"] = s;
      std::string mig = EncodeMigAllocations(it->second);
      if (!mig.empty()) annos["hami.io/vgpu-mig-allocations"] = mig;
    }
  }
};

// parseMetaxAnnos, metax/device.go:168-183: json.Unmarshal into map[int]int, then the entry at `index`; ANY error
// (syntax, a key that is not an integer, a value that is not an integer number) and a missing key give 0.
// Restated for the documents the annotation holds: one object of "int": int|null pairs (or the document null); string
// escapes inside keys are not decoded (such a key counts as an error).
static bool ParseInt64(const std::string& v, long long& out) {   // strconv.ParseInt(v, 10, 64)
  size_t i = 0; bool neg = false;
  if (!v.empty() && (v[0] == '+' || v[0] == '-')) { neg = v[0] == '-'; i = 1; }
  if (i >= v.size()) return false;
  unsigned long long acc = 0;
  for (; i < v.size(); i++) {
    if (v[i] < '0' || v[i] > '9') return false;
    unsigned long long d = (unsigned long long)(v[i] - '0');
    if (acc > (0x8000000000000000ULL - d) / 10) return false;
    acc = acc * 10 + d;
  }
  if (!neg && acc > 0x7fffffffffffffffULL) return false;
  out = neg ? (long long)(0 - acc) : (long long)acc;
  return true;
}

float parseMetaxAnnos(const std::string& s, int index) {
  size_t i = 0;
  auto ws = [&] { while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) i++; };
  std::map<long long, long long> m;
  ws();
  if (s.compare(i, 4, "null") == 0) { i += 4; ws(); return 0; }
  if (i >= s.size() || s[i] != '{') return 0;
  i++; ws();
  if (i < s.size() && s[i] == '}') { i++; ws(); return 0; }
  while (true) {
    if (i >= s.size() || s[i] != '"') return 0;
    size_t e = s.find('"', i + 1);
    if (e == std::string::npos) return 0;
    std::string key = s.substr(i + 1, e - i - 1);
    if (key.find('\\') != std::string::npos) return 0;
    long long kv;
    if (!ParseInt64(key, kv)) return 0;
    i = e + 1; ws();
    if (i >= s.size() || s[i] != ':') return 0;
    i++; ws();
    if (s.compare(i, 4, "null") == 0) { i += 4; m[kv] = 0; }   // a fresh zero element is stored (encoding/json decode.go object())
    else {
      size_t b = i;
      if (i < s.size() && s[i] == '-') i++;
      if (i >= s.size() || s[i] < '0' || s[i] > '9') return 0;
      if (s[i] == '0' && i + 1 < s.size() && s[i + 1] >= '0' && s[i + 1] <= '9') return 0;   // JSON: no leading zeros
      while (i < s.size() && s[i] >= '0' && s[i] <= '9') i++;
      if (i < s.size() && (s[i] == '.' || s[i] == 'e' || s[i] == 'E')) return 0;            // not an integer: UnmarshalTypeError
      long long vv;
      if (!ParseInt64(s.substr(b, i - b), vv)) return 0;
      m[kv] = vv;
    }
    ws();
    if (i >= s.size()) return 0;
    if (s[i] == ',') { i++; ws(); continue; }
    if (s[i] == '}') { i++; break; }
    return 0;
  }
  ws();
  if (i != s.size()) return 0;
  auto it = m.find(index);
  return it == m.end() ? 0.0f : (float)it->second;
}

// ---- AWS Neuron (pkg/device/awsneuron/device.go) --------------------------------------------------------------
// Totalcore / Usedcores are core BIT MASKS; CustomInfo["awsusageinfo"] (DeviceUsage.AwsUsage here) is the mask the pod's
// earlier containers already hold on the device.  CustomInfo["AWSNodeType"] is HasNetworkID / NetworkID (see oracle.h).
int32_t countMaskAvailable(int32_t mask) {          // awsneuron/device.go:357-365
  int32_t ret = 0;
  for (int32_t tmp = mask; tmp > 0; tmp /= 2) ret += tmp % 2;
  return ret;
}
int addCoreUsage(int prev /* -1 = key absent */, int require) {   // :367-389
  const int count = prev < 0 ? 0 : prev;
  if (count == 0) {
    if (require == 2) return 3;
    if (require == 1) return 1;
  }
  return 3 - count;
}
std::vector<int> continuousDeviceAvailable(const std::vector<DeviceUsage*>& devices, int start, int count) {   // :390-404
  if ((int)devices.size() < start + count) return {};
  std::vector<int> res;
  for (int it = start; it < start + count; it++) {
    if (devices[it]->Used > 0 || !devices[it]->Health) return {};
    res.push_back(it);
  }
  return res;
}
std::vector<int> graphSelect(const std::vector<DeviceUsage*>& devices, int count) {   // :406-441
  if (devices.empty() || !devices[0]->HasNetworkID) return {};
  if (devices[0]->NetworkID == 1) {               // instance type contains "inf" / "Inf": a ring, any start
    for (int start = 0; start < (int)devices.size(); start++) {
      auto res = continuousDeviceAvailable(devices, start, count);
      if (!res.empty()) return res;
    }
    return {};
  }
  if (count == 1 || count == 4 || count == 8 || count == 16) {
    for (int start = 0; start < (int)devices.size(); start += count) {
      auto res = continuousDeviceAvailable(devices, start, count);
      if (!res.empty()) return res;
    }
  }
  return {};
}
FitResult fitNeuron(const FitGates& g, const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request,
                    const Pod& pod, const std::string& gpuPolicyDefault) {   // :443-541
  const ContainerDeviceRequest& k = request;
  std::map<std::string, int> reason;
  ContainerDevices tmpDevs;
  const bool isMutex = PolicyContains(GetGPUSchedulerPolicyByPod(gpuPolicyDefault, pod), "mutex");
  if (k.Nums > 1) {
    std::vector<int> alloc = graphSelect(devices, (int)request.Nums);
    if (alloc.empty()) {
      reason["NumaNotFit"]++;
      FitResult r; r.fit = false; r.reason = GenReason(reason, (int)devices.size());
      return r;
    }
    for (int dev : alloc)
      for (auto val : devices)
        if (val->Index == (unsigned)dev) {
          ContainerDevice cd;
          cd.Idx = (int)val->Index; cd.UUID = val->ID; cd.Type = request.Type;
          cd.Usedmem = val->Totalmem; cd.Usedcores = val->Totalcore;
          cd.AwsUsage = addCoreUsage(val->AwsUsage, (int)k.Coresreq);
          tmpDevs.push_back(cd);
          break;
        }
    FitResult r; r.fit = true; r.devs = tmpDevs;
    return r;
  }
  for (int i = (int)devices.size() - 1; i >= 0; i--) {
    DeviceUsage* dev = devices[i];
    if (dev->AwsUsage < 0) dev->AwsUsage = (int)dev->Usedcores;        // written before any gate, also on unhealthy devices
    if (!dev->Health) { reason["CardNotHealth"]++; continue; }
    if (!Contains(dev->Type, k.Type)) { reason["CardTypeMismatch"]++; continue; }
    if (k.Type != g.commonWord) { reason["CardTypeMismatch"]++; continue; }   // checkType, :268-273
    if (!CheckUUID(pod.Annotations, dev->ID, g.useUUID, g.noUseUUID)) { reason["CardUuidMismatch"]++; continue; }
    if (dev->Count <= dev->Used) { reason["CardTimeSlicingExhausted"]++; continue; }
    if (isMutex && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
    if (i32sub(countMaskAvailable(dev->Totalcore), countMaskAvailable(dev->Usedcores)) < k.Coresreq) { reason["CardInsufficientCore"]++; continue; }
    ContainerDevice cd;
    cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = k.Type; cd.Usedmem = 0;
    cd.AwsUsage = addCoreUsage(dev->AwsUsage, (int)k.Coresreq);
    cd.Usedcores = (int32_t)cd.AwsUsage;
    tmpDevs.push_back(cd);
    FitResult r; r.fit = true; r.devs = tmpDevs;
    return r;
  }
  FitResult r; r.fit = false; r.reason = GenReason(reason, (int)devices.size());
  return r;
}

// ---- kunlun (whole XPUs): kunlun/topo.go + kunlun/device.go:184-231 ---------------------------------------------
// Fit runs graghSelect over the typeDevices in LIST order (wings = positions 0-3 / 4-7), FitXPU = idle and healthy,
// the two use/no-use uuid annotation pairs applied inside fitFn; a failure reads the number of devices a uuid list
// turned away, else one NumaNotFit.
// vxpu = the shared "XPU" device type (kunlun/vdevice.go:231-306): the same graghSelect with FitVXPU — healthy, the
// request's memory still fits, and a card in use only takes requests of the size it already serves (its average) —
// idle cards only under the mutex policy; the pick records the request's memory and cores.
static bool fitVXPU(const DeviceUsage& d, const ContainerDeviceRequest& request) {          // vdevice.go:293-306
  if (!d.Health) return false;
  if (i32add(request.Memreq, d.Usedmem) > d.Totalmem) return false;
  if (d.Used == 0) return true;
  return d.Usedmem / d.Used == request.Memreq;
}
FitResult fitKunlun(const std::string& word, const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                    bool vxpu, const std::string& gpuPolicyDefault) {
  const bool isMutex = vxpu && PolicyContains(GetGPUSchedulerPolicyByPod(gpuPolicyDefault, pod), "mutex");
  std::set<std::string> uuidMismatches;
  std::vector<bool> fit;
  std::vector<int> index;
  for (auto d : devices) {
    bool ok;
    if (!CheckUUID(pod.Annotations, d->ID, "hami.io/use-xpu-uuid", "hami.io/no-use-xpu-uuid") ||
        !CheckUUID(pod.Annotations, d->ID, "baidu.com/use-gpuuuid", "baidu.com/nouse-gpuuuid")) {
      uuidMismatches.insert(d->ID);
      ok = false;
    } else if (vxpu) ok = (!isMutex || d->Used == 0) && fitVXPU(*d, request);
    else ok = d->Used == 0 && d->Health;            // FitXPU, device.go:238-240
    fit.push_back(ok);
    index.push_back((int)d->Index);
  }
  std::vector<int> alloc = KunlunGraphSelect(fit, index, (int)request.Nums);
  FitResult r;
  if (alloc.empty()) {
    std::map<std::string, int> reason;
    if (isMutex)                                     // vdevice.go:259-266
      for (auto d : devices) if (d->Used > 0 && fitVXPU(*d, request)) reason["ExclusiveDeviceAllocateConflict"]++;
    if (reason.empty() && !uuidMismatches.empty()) reason["CardUuidMismatch"] += (int)uuidMismatches.size();
    if (reason.empty()) reason["NumaNotFit"]++;
    r.fit = false; r.reason = GenReason(reason, (int)devices.size());
    return r;
  }
  for (int dev : alloc)
    for (auto val : devices)
      if (val->Index == (unsigned)dev) {
        ContainerDevice cd;
        cd.Idx = (int)val->Index; cd.UUID = val->ID; cd.Type = request.Type;
        cd.Usedmem = vxpu ? request.Memreq : val->Totalmem;
        cd.Usedcores = vxpu ? request.Coresreq : val->Totalcore;
        r.devs.push_back(cd);
        break;
      }
  r.fit = true;
  (void)word;
  return r;
}

// ---- enflame shared vGCU ("Enflame", enflame/device.go:382-478): one DRS profile per request, chosen from the profiles
// the node's devices advertise, then the template gates with the profile's memory and core share --------------------
struct DrsProfile { std::string Name, ID; long long Size = 0, MemoryGB = 0, CorePercent = 0; int RequestMemoryGB = 0, RequestCorePercent = 0; };
static void parseDrsProfile(const std::string& profileName, long long& size, long long& memGB) {     // parseProfile, :621-643 (Go int is 64 bit)
  size = 0; memGB = 0;
  std::string normalized = ToLower(TrimSpace(profileName));
  std::vector<std::string> parts = Split(normalized, '.');
  if (parts.size() < 2) return;
  std::string sizePart = TrimSpace(parts[0]);
  if (sizePart.size() >= 1 && sizePart.back() == 'g') sizePart.pop_back();
  long long sz = 0;
  if (!ParseInt64(sizePart, sz)) return;                         // strconv.Atoi
  std::string memPart = TrimSpace(parts[1]);
  if (memPart.size() >= 2 && memPart.compare(memPart.size() - 2, 2, "gb") == 0) memPart.erase(memPart.size() - 2);
  long long mg = 0;
  size = sz;
  if (!ParseInt64(memPart, mg)) return;
  memGB = mg;
}
// collectDRSProfiles, :567-619.  Candidates with equal (Size, MemoryGB) come out of a Go map in random order and an
// unstable sort: taken in ascending profile name (canonical rule, like the other map-order cases).
static std::vector<DrsProfile> collectDrsProfiles(const std::vector<DeviceUsage*>& devices) {
  long long maxSlice = 0;
  for (auto d : devices) if (d->EnfMaxSlice > maxSlice) maxSlice = d->EnfMaxSlice;
  if (maxSlice <= 0)
    for (auto d : devices)
      for (auto& kv : d->EnfProfiles) { long long size, mem; parseDrsProfile(kv.first, size, mem); if (size > maxSlice) maxSlice = size; }
  if (maxSlice <= 0) return {};
  std::map<std::string, DrsProfile> seen;
  for (auto d : devices)
    for (auto& kv : d->EnfProfiles) {
      long long size, memGB;
      parseDrsProfile(kv.first, size, memGB);
      if (size <= 0 || memGB <= 0) continue;
      DrsProfile c;
      c.Name = kv.first; c.ID = kv.second; c.Size = size; c.MemoryGB = memGB;
      c.CorePercent = (long long)std::ceil((double)size * 100 / (double)maxSlice);
      seen[kv.first] = c;
    }
  std::vector<DrsProfile> candidates;
  for (auto& kv : seen) candidates.push_back(kv.second);          // ascending name
  std::stable_sort(candidates.begin(), candidates.end(), [](const DrsProfile& a, const DrsProfile& b) {
    if (a.Size == b.Size) return a.MemoryGB < b.MemoryGB;
    return a.Size < b.Size;
  });
  return candidates;
}
static bool selectDrsProfile(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, DrsProfile& out) {   // :498-531
  std::vector<DrsProfile> candidates = collectDrsProfiles(devices);
  if (candidates.empty()) return false;
  if (request.MemPercentagereq == 0) {                              // enflameRequestModeDirect: the slice count itself
    for (auto& c : candidates) if (c.Size == (long long)request.Memreq) { out = c; return true; }
    return false;
  }
  const long long maxMemGB = candidates.back().MemoryGB;
  int requestMemGB;                                                 // normalizeMemoryRequestToGB, :645-657
  if (request.Memreq <= 0) requestMemGB = 0;
  else if (maxMemGB <= 0) requestMemGB = (int)request.Memreq;
  else if ((long long)request.Memreq > (long long)maxMemGB * 2) requestMemGB = (int)std::ceil((double)request.Memreq / 1024.0);
  else requestMemGB = (int)request.Memreq;
  const int requestCorePercent = (int)request.Coresreq;
  for (auto& c : candidates) {
    if (requestMemGB > 0 && c.MemoryGB < requestMemGB) continue;
    if (requestCorePercent > 0 && c.CorePercent < requestCorePercent) continue;
    out = c;
    out.RequestMemoryGB = requestMemGB; out.RequestCorePercent = requestCorePercent;
    return true;
  }
  return false;
}
FitResult fitEnflame(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                     const std::string& gpuPolicyDefault) {
  ContainerDeviceRequest k = request;
  const int32_t originReq = k.Nums;
  std::map<std::string, int> reason;
  const bool isMutex = PolicyContains(GetGPUSchedulerPolicyByPod(gpuPolicyDefault, pod), "mutex");
  FitResult r;
  DrsProfile profile;
  auto modeNotFit = [&]() { reason["ModeNotFit"]++; r.fit = false; r.reason = GenReason(reason, (int)devices.size()); return r; };
  if (!selectDrsProfile(devices, k, profile)) return modeNotFit();
  if (profile.Size <= 0 || profile.Size > 2147483647LL) return modeNotFit();
  if (profile.MemoryGB <= 0 || profile.MemoryGB > 2147483647LL / 1024) return modeNotFit();
  if (profile.CorePercent > 2147483647LL) return modeNotFit();
  const int32_t profileMemoryMiB = (int32_t)(profile.MemoryGB * 1024);
  const int32_t profileCorePercent = (int32_t)profile.CorePercent;
  for (int i = (int)devices.size() - 1; i >= 0; i--) {
    DeviceUsage* dev = devices[i];
    if (!dev->Health) { reason["CardNotHealth"]++; continue; }
    if (k.Type != "Enflame") { reason["CardTypeMismatch"]++; continue; }                       // checkType, :310-315
    if (!CheckUUID(pod.Annotations, dev->ID, "enflame.com/use-gpuuuid", "enflame.com/nouse-gpuuuid")) { reason["CardUuidMismatch"]++; continue; }
    if (dev->Count <= dev->Used) { reason["CardTimeSlicingExhausted"]++; continue; }
    if (isMutex && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
    if (i32sub(dev->Totalmem, dev->Usedmem) < profileMemoryMiB) { reason["CardInsufficientMemory"]++; continue; }
    if (dev->Totalcore > 0 && i32sub(dev->Totalcore, dev->Usedcores) < profileCorePercent) { reason["CardInsufficientCore"]++; continue; }
    if (k.Nums > 0) {
      k.Nums--;
      ContainerDevice cd;
      cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = k.Type; cd.Usedmem = profileMemoryMiB; cd.Usedcores = profileCorePercent;
      cd.HasDrs = true; cd.DrsProfileName = profile.Name; cd.DrsProfileID = profile.ID; cd.DrsMinor = dev->EnfMinor; cd.DrsIndex = dev->EnfIndex;
      cd.DrsSlice = (int)profile.Size; cd.DrsRequestMem = profile.RequestMemoryGB; cd.DrsRequestCore = profile.RequestCorePercent;
      r.devs.push_back(cd);
    }
    if (k.Nums == 0) { r.fit = true; return r; }
  }
  if (!r.devs.empty()) reason["AllocatedCardsInsufficientRequest"] = (int)r.devs.size();
  r.fit = false; r.reason = GenReason(reason, (int)devices.size());
  (void)originReq;
  return r;
}

// ---- metax sGPU ("Metax-SGPU", metax/sdevice.go:332-495): filter, then prioritise ------------------------------------
struct SgpuState {                      // the vendor object's jqCache (metax/qos.go): device UUID -> the QoS it is about to switch to
  std::map<std::string, std::string> jq;
};
static const std::string* podAnno(const Pod& pod, const char* key) { return anno(pod, key); }
static bool sgpuCheckAppClass(const std::string& requestClass, const DeviceUsage& dev) {    // checkAppClass, sdevice.go:738-775
  if (dev.Used == 0) return true;
  const bool devOnline = dev.OnlinePods > 0, devOffline = dev.OfflinePods > 0;
  if ((devOffline || devOnline) == (requestClass == "offline" || requestClass == "online")) {
    if (devOnline && requestClass == "offline") return false;
    return true;
  }
  return false;
}
static bool sgpuCheckQos(const SgpuState& st, const std::string& reqQos, const DeviceUsage& usage, const ContainerDeviceRequest& request) {   // :527-556
  if (usage.Used == 0) return true;
  if (request.Coresreq == 100) return true;
  std::string devQos;
  auto it = st.jq.find(usage.ID);
  if (it != st.jq.end()) devQos = it->second;
  else if (usage.HasSgpu) devQos = usage.QosPolicy;
  return devQos == "" || reqQos == devQos;
}
// prioritizeOnlineDevices, :717-770: sort.Slice on <= 12 candidates is Go's insertion sort
static std::vector<DeviceUsage*> sgpuPrioritizeOnline(std::vector<DeviceUsage*> c, int require) {
  if ((int)c.size() <= require) return c;
  auto less = [](const DeviceUsage* a, const DeviceUsage* b) {
    const bool iOnly = a->OnlinePods == 0 && a->OfflinePods > 0, jOnly = b->OnlinePods == 0 && b->OfflinePods > 0;
    if (iOnly && jOnly) return a->OfflinePods < b->OfflinePods;
    if (iOnly) return false;
    if (jOnly) return true;
    if (a->OnlinePods == b->OnlinePods) return a->OfflinePods < b->OfflinePods;
    return a->OnlinePods < b->OnlinePods;
  };
  for (size_t i = 1; i < c.size(); i++)
    for (size_t j = i; j > 0 && less(c[j], c[j - 1]); j--) std::swap(c[j], c[j - 1]);
  c.resize(require);
  return c;
}
// prioritizeExclusiveDevices, :583-645.  The MetaLink groups come out of a Go map: taken in ascending link zone (canonical
// rule), then ordered by size with Go's insertion sort.
static std::vector<DeviceUsage*> sgpuPrioritizeExclusive(const std::vector<DeviceUsage*>& cand, int require) {
  if ((int)cand.size() <= require) return cand;
  std::map<int32_t, std::vector<DeviceUsage*>> byZone;
  for (auto d : cand) byZone[d->HasSgpu ? d->LinkZone : 0].push_back(d);
  std::vector<std::vector<DeviceUsage*>> link;
  std::vector<DeviceUsage*> other;
  for (auto& kv : byZone) {
    if (kv.first <= 0) other.insert(other.end(), kv.second.begin(), kv.second.end());
    else link.push_back(kv.second);
  }
  for (size_t i = 1; i < link.size(); i++)
    for (size_t j = i; j > 0 && link[j].size() < link[j - 1].size(); j--) std::swap(link[j], link[j - 1]);
  for (auto& g : link)
    if ((int)g.size() >= require) return std::vector<DeviceUsage*>(g.begin(), g.begin() + require);
  std::vector<DeviceUsage*> out;
  for (auto& g : link)
    for (auto d : g) { out.push_back(d); if ((int)out.size() >= require) return out; }
  for (auto d : other) { out.push_back(d); if ((int)out.size() >= require) return out; }
  return std::vector<DeviceUsage*>(cand.begin(), cand.begin() + require);
}
FitResult fitSgpu(const SgpuState& st, const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                  const std::string& gpuPolicyDefault) {
  std::map<std::string, int> reason;
  const bool isMutex = PolicyContains(GetGPUSchedulerPolicyByPod(gpuPolicyDefault, pod), "mutex");
  const std::string* classAnno = podAnno(pod, "metax-tech.com/sgpu-app-class");
  const std::string appClass = classAnno ? *classAnno : "";
  const std::string* qosAnno = podAnno(pod, "metax-tech.com/sgpu-qos-policy");
  const std::string reqQos = qosAnno ? *qosAnno : "";
  std::vector<DeviceUsage*> candidates;
  for (int i = (int)devices.size() - 1; i >= 0; i--) {
    DeviceUsage* dev = devices[i];
    if (request.Type != "Metax-SGPU") { reason["CardTypeMismatch"]++; continue; }
    if (!dev->Health) { reason["CardUnhealthy"]++; continue; }
    if (!sgpuCheckAppClass(appClass, *dev)) { reason["CardAppClassMismatch"]++; continue; }
    if (appClass != "online" && !sgpuCheckQos(st, reqQos, *dev, request)) { reason["CardQosPolicyMismatch"]++; continue; }
    if (!CheckUUID(pod.Annotations, dev->ID, "metax-tech.com/use-gpuuuid", "metax-tech.com/nouse-gpuuuid")) { reason["CardUuidMismatch"]++; continue; }
    if (dev->Count <= dev->Used) { reason["CardTimeSlicingExhausted"]++; continue; }
    if (isMutex && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
    int32_t memreq = request.Memreq > 0 ? request.Memreq : i32mul(dev->Totalmem, request.MemPercentagereq) / 100;
    if (i32sub(dev->Totalmem, dev->Usedmem) < memreq) { reason["CardInsufficientMemory"]++; continue; }
    if (appClass != "online") {
      if (i32sub(dev->Totalcore, dev->Usedcores) < request.Coresreq) { reason["CardInsufficientCore"]++; continue; }
      if (dev->Totalcore == 100 && request.Coresreq == 100 && dev->Used > 0) { reason["ExclusiveDeviceAllocateConflict"]++; continue; }
      if (dev->Totalcore != 0 && dev->Usedcores == dev->Totalcore && request.Coresreq == 0) { reason["CardComputeUnitsExhausted"]++; continue; }
    }
    candidates.push_back(dev);
  }
  FitResult r;
  if ((int)candidates.size() < (int)request.Nums) {
    if (!candidates.empty()) reason["AllocatedCardsInsufficientRequest"] = (int)candidates.size();
    r.fit = false; r.reason = GenReason(reason, (int)devices.size());
    return r;
  }
  const int nums = request.Nums < 0 ? 0 : (int)request.Nums;
  std::vector<DeviceUsage*> best(candidates.begin(), candidates.begin() + nums);
  if ((int)candidates.size() > nums) {
    if (classAnno) { if (appClass == "online") best = sgpuPrioritizeOnline(candidates, nums); }
    else if (request.Coresreq == 100) best = sgpuPrioritizeExclusive(candidates, nums);
  }
  const int32_t coreReq = appClass == "online" ? 0 : request.Coresreq;
  for (auto dev : best) {
    ContainerDevice cd;
    cd.Idx = (int)dev->Index; cd.UUID = dev->ID; cd.Type = dev->Type;
    cd.Usedmem = request.Memreq > 0 ? request.Memreq : i32mul(dev->Totalmem, request.MemPercentagereq) / 100;
    cd.Usedcores = coreReq;
    cd.HasSgpu = dev->HasSgpu; cd.SgpuQos = dev->QosPolicy; cd.SgpuLinkZone = dev->LinkZone; cd.PodAnnotations = &pod.Annotations;
    r.devs.push_back(cd);
  }
  r.fit = true;
  return r;
}
// LinkDevices.Score, metax/score.go:27-57
static int sgpuLinkScore(const std::vector<std::pair<std::string, int32_t>>& devs) {
  int score = 0;
  for (size_t i = 0; i < devs.size(); i++)
    for (size_t j = i + 1; j < devs.size(); j++)
      if (devs[i].first != devs[j].first && devs[i].second != 0 && devs[j].second != 0 && devs[i].second == devs[j].second) score += 10;
  return score;
}
static float sgpuScoreNode(bool topoDefault, const PodSingleDevice& podDevices, const std::vector<DeviceUsage>* previous) {   // sdevice.go:295-304
  static const std::vector<DeviceUsage> none;
  const std::vector<DeviceUsage>& prev = previous ? *previous : none;
  bool online = false;                                              // appClassOnlineEnable, :705-715
  for (auto& ctr : podDevices)
    for (auto& d : ctr)
      if (d.PodAnnotations) { auto it = d.PodAnnotations->find("metax-tech.com/sgpu-app-class"); if (it != d.PodAnnotations->end() && it->second == "online") online = true; }
  if (online) {                                                     // scoreOnlineDevices, :772-811
    std::set<std::string> deviceSet;
    for (auto& ctr : podDevices) for (auto& d : ctr) deviceSet.insert(d.UUID);
    int result = 0;
    for (auto& dev : prev) {
      if (!deviceSet.count(dev.ID)) continue;
      if (dev.OfflinePods > 0 && dev.OnlinePods == 0) result += -100 * dev.OfflinePods;
      else result += -20 * dev.OnlinePods - 2 * dev.OfflinePods;
    }
    return (float)result;
  }
  bool topo = false, decided = false;                               // topologyAwareEnable, :647-667: the FIRST exclusive pick decides
  for (auto& ctr : podDevices) {
    for (auto& d : ctr) {
      if (d.Usedcores != 100) continue;
      topo = topoDefault;
      if (d.PodAnnotations) {
        auto it = d.PodAnnotations->find("metax-tech.com/sgpu-topology-aware");
        bool b;
        if (it != d.PodAnnotations->end() && ParseBool(it->second, b)) topo = b;
      }
      decided = true;
      break;
    }
    if (decided) break;
  }
  if (!topo) return 0;
  std::vector<std::pair<std::string, int32_t>> available, allocated, rest;     // scoreExclusiveDevices
  for (auto& ctr : podDevices)
    for (auto& d : ctr)
      if (d.Usedcores == 100) allocated.emplace_back(d.UUID, d.HasSgpu ? d.SgpuLinkZone : 0);
  for (auto& dev : prev) {
    if (dev.Used != 0) continue;
    const int32_t zone = dev.HasSgpu ? dev.LinkZone : 0;
    available.emplace_back(dev.ID, zone);
    bool find = false;
    for (auto& a : allocated) if (a.first == dev.ID) { find = true; break; }
    if (!find) rest.emplace_back(dev.ID, zone);
  }
  const int availableScore = sgpuLinkScore(available), allocatedScore = sgpuLinkScore(allocated), restScore = sgpuLinkScore(rest);
  const int lossScore = availableScore - allocatedScore - restScore;
  return (float)(10 * allocatedScore - lossScore);
}

// iluvatar (one instance per CommonWord, iluvatar/device.go:60-83) and biren (biren/device.go:54-66)
class PlainDevices : public Devices {
 public:
  explicit PlainDevices(const Config::PlainVendor& v) : v_(v) {}
  std::string CommonWord() const override { return v_.CommonWord; }
  FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                const NodeInfo* nodeInfo, const std::string& def, const PodDevices* allocated,
                const QuotaManager* quota) const override {
    static const Vendor kinds[] = {Vendor::PLAIN, Vendor::COUNT_ONLY, Vendor::AMD, Vendor::MTHREADS, Vendor::WHOLE_CARD, Vendor::METAX,
                                   Vendor::PLAIN /* 6 = Neuron, below */, Vendor::VASTAI};
    if (v_.FitKind == 11) return fitSgpu(sgpu_, devices, request, pod, def);
    if (v_.FitKind == 10) return fitEnflame(devices, request, pod, def);
    if (v_.FitKind == 8 || v_.FitKind == 9) return fitKunlun(v_.CommonWord, devices, request, pod, v_.FitKind == 9, def);
    FitGates g{kinds[v_.FitKind >= 0 && v_.FitKind <= 7 ? v_.FitKind : 0],
               v_.CommonWord,
               v_.UseType, v_.NoUseType, v_.UseUUID, v_.NoUseUUID};
    if (v_.FitKind == 6) return fitNeuron(g, devices, request, pod, def);
    return fitTemplate(g, devices, request, pod, nodeInfo, def, allocated, quota, GetResourceNames());
  }
  bool AddResourceUsage(DeviceUsage& n, ContainerDevice& ctr) const override {
    if (v_.FitKind == 11) {                                                // metax/sdevice.go:310-330
      Devices::AddResourceUsage(n, ctr);
      const std::string* cls = ctr.PodAnnotations ? (ctr.PodAnnotations->count("metax-tech.com/sgpu-app-class") ? &ctr.PodAnnotations->at("metax-tech.com/sgpu-app-class") : nullptr) : nullptr;
      if (!(cls && *cls == "online") && n.HasSgpu) {
        std::string expectedQos;
        if (ctr.PodAnnotations) { auto q = ctr.PodAnnotations->find("metax-tech.com/sgpu-qos-policy"); if (q != ctr.PodAnnotations->end()) expectedQos = q->second; }
        if (ctr.Usedcores == 100) expectedQos = "";
        n.QosPolicy = expectedQos;        // the working copy's CustomInfo["QosPolicy"]
      }
      return true;
    }
    if (v_.FitKind == 8) { n.Used = i32add(n.Used, 1); return true; }      // kunlun/device.go:179-182: the slot only
    if (v_.FitKind == 10) {                                                // enflame/device.go:372-380: the slot count grows by the slice size
      int slice = ctr.DrsSlice <= 0 ? 1 : ctr.DrsSlice;
      long long u = (long long)n.Used + slice;
      n.Used = (int32_t)std::min<long long>(std::max<long long>(u, -2147483648LL), 2147483647LL);      // clampToInt32
      n.Usedcores = i32add(n.Usedcores, ctr.Usedcores);
      n.Usedmem = i32add(n.Usedmem, ctr.Usedmem);
      return true;
    }
    Devices::AddResourceUsage(n, ctr);
    if (v_.FitKind == 6) {                       // awsneuron/device.go:340-355
      if (n.AwsUsage < 0) n.AwsUsage = 0;
      if (ctr.AwsUsage >= 0) n.AwsUsage += ctr.AwsUsage;
    }
    return true;
  }
  // metax/device.go:185-214: 2000 minus the node's topology loss (binpack) / score (spread) for this many devices
  float ScoreNode(const PodSingleDevice& podDevices, const std::string& policy, const NodeInfo* nodeInfo,
                  const std::vector<DeviceUsage>* previous) const override {
    if (v_.FitKind == 11) return sgpuScoreNode(v_.Option != 0, podDevices, previous);
    if (v_.FitKind == 8) {            // kunlun/device.go:154-177
      std::vector<int> current, prev;
      auto addidx = [](std::vector<int>& t, int v) { if (std::find(t.begin(), t.end(), v) == t.end()) t.push_back(v); };
      if (previous)
        for (auto& d : *previous) {
          if (!Contains(d.Type, "kunlun")) return 0;
          if (d.Used > 0) addidx(prev, (int)d.Index);
        }
      for (auto& ctr : podDevices)
        for (auto& val : ctr) {
          if (!Contains(val.Type, "kunlun")) return 0;
          addidx(current, val.Idx);
        }
      return KunlunCalcScore(prev, current);
    }
    if (v_.FitKind == 7) {            // vastai/device.go:181-216: largest share of one card among a container's dies, unweighted
      float score = 0;
      for (auto& containerDevices : podDevices) {
        if (containerDevices.empty()) continue;
        std::map<int, int> cntMap;
        for (auto& d : containerDevices) {
          if (!d.HasNetworkID) return 0;          // no CustomInfo / not a die / no AIC
          cntMap[d.NetworkID]++;
        }
        int maxCnt = 0, totalCnt = 0;
        for (auto& kv : cntMap) { if (kv.second > maxCnt) maxCnt = kv.second; totalCnt += kv.second; }
        if (totalCnt == 0) continue;
        score = score + (float)maxCnt / (float)totalCnt;
      }
      return score;
    }
    if (v_.FitKind != 5) return 0;
    int sum = 0;
    for (auto& ctr : podDevices) sum += (int)ctr.size();
    float res = 0;
    if (policy == "binpack") {
      if (nodeInfo && nodeInfo->MetaxLosses) res = (float)2000 - parseMetaxAnnos(*nodeInfo->MetaxLosses, sum);
    } else if (policy == "spread") {
      if (nodeInfo && nodeInfo->MetaxScores) res = (float)2000 - parseMetaxAnnos(*nodeInfo->MetaxScores, sum);
    }
    return res;
  }
  void PatchAnnotations(const Pod& pod, std::map<std::string, std::string>& annos, const PodDevices& pd) const override {
    // iluvatar/device.go:143-154, biren/device.go:167-177, metax/device.go:91-103
    auto it = pd.find(v_.CommonWord);
    if (it != pd.end() && !it->second.empty()) {
      std::string s = EncodePodSingleDevice(it->second);
      if (v_.FitKind == 6) {
        // awsneuron/device.go:200-243: the allocated key only.  AWS_NEURON_IDS / NEURON_RESOURCE_TYPE / NEURON_ALLOC_TIME /
        // NEURON_ALLOCATED / predicate-node depend on the container's resource LIMITS (asked by device or by core), which
        // the path's input does not carry: the caller's own PatchAnnotations adds them from the returned PodDevices.
        annos[v_.Allocated] = s;
        return;
      }
      if (v_.FitKind == 3) {
        // mthreads/device.go:155-177: only the allocated key; the card indices of the LAST container that got devices;
        // the node again under its own key; a timestamp (blanked, canonical rule 6)
        annos[v_.Allocated] = s;
        for (auto& ctr : it->second) {
          if (ctr.empty()) continue;
          std::string value;
          for (auto& d : ctr) { if (!value.empty()) value += ","; value += std::to_string(d.Idx); }
          annos["mthreads.com/gpu-index"] = value;
          annos["mthreads.com/predicate-time"] = "";
          annos["mthreads.com/predicate-node"] = annos["hami.io/vgpu-node"];
        }
        return;
      }
      if (v_.FitKind == 11) {
        // metax/sdevice.go:142-171: both hami keys, the pod's devices in metax's own JSON, a timestamp (blanked), and the
        // side effect that makes the next Filter see a QoS switch in flight (addJitteryQos, :558-581)
        annos[v_.ToAllocate] = s;
        annos[v_.Allocated] = s;
        std::string js = "[";
        for (size_t c = 0; c < it->second.size(); c++) {
          js += c ? ",[" : "[";
          for (size_t d = 0; d < it->second[c].size(); d++) {
            const ContainerDevice& cd = it->second[c][d];
            js += d ? ",{" : "{";
            js += "\"uuid\":" + jsonString(cd.UUID);
            if (cd.Usedcores != 0) js += ",\"compute\":" + std::to_string(cd.Usedcores);
            if (cd.Usedmem != 0) js += ",\"vRam\":" + std::to_string(cd.Usedmem);
            js += "}";
          }
          js += "]";
        }
        annos["metax-tech.com/gpu-devices-allocated"] = js + "]";
        annos["metax-tech.com/predicate-time"] = "";
        const std::string* cls = anno(pod, "metax-tech.com/sgpu-app-class");
        if (!(cls && *cls == "online")) {
          const std::string* q = anno(pod, "metax-tech.com/sgpu-qos-policy");
          const std::string reqQos = q ? *q : "";
          for (auto& ctrdev : it->second)
            for (auto& dev : ctrdev) {
              if (!dev.HasSgpu) continue;
              const std::string expectedQos = dev.Usedcores == 100 ? "" : reqQos;
              if (dev.SgpuQos != expectedQos) sgpu_.jq[dev.UUID] = expectedQos;
            }
        }
        return;
      }
      if (v_.FitKind == 10) {
        // enflame/device.go:254-300: the allocated key, gcu-assigned = "false", a timestamp (blanked, canonical rule 6), the
        // first container's index / minor / slice count, and assigned-containers = {container name: request} as JSON
        annos[v_.Allocated] = s;
        annos["enflame.com/gcu-assigned"] = "false";
        annos["enflame.com/gcu-assigned-time"] = "";
        std::map<std::string, std::string> assigned;
        for (size_t ctridx = 0; ctridx < it->second.size(); ctridx++) {
          const ContainerDevices& ctrDevices = it->second[ctridx];
          if (ctrDevices.empty()) continue;
          const ContainerDevice& chosen = ctrDevices[0];
          int slice = chosen.DrsSlice <= 0 ? 1 : chosen.DrsSlice;
          std::string ctrName = ctridx < pod.ContainerNames.size() ? pod.ContainerNames[ctridx] : "container-" + std::to_string(ctridx);
          std::string o = "{\"allocated\":false,\"request\":" + std::to_string(slice);
          if (!chosen.DrsProfileID.empty()) o += ",\"profileID\":" + jsonString(chosen.DrsProfileID);
          if (!chosen.DrsProfileName.empty()) o += ",\"profileName\":" + jsonString(chosen.DrsProfileName);
          assigned[ctrName] = o + "}";
          if (!annos.count("enflame.com/gcu-assigned-index") && !chosen.DrsIndex.empty()) {
            annos["enflame.com/gcu-assigned-index"] = chosen.DrsIndex;
            annos["enflame.com/gcu-assigned-id"] = chosen.DrsIndex;
          }
          if (!annos.count("enflame.com/gcu-assigned-minor") && !chosen.DrsMinor.empty()) annos["enflame.com/gcu-assigned-minor"] = chosen.DrsMinor;
          if (!annos.count("enflame.com/gcu-request-size") && slice > 0) annos["enflame.com/gcu-request-size"] = std::to_string(slice);
        }
        if (!assigned.empty()) {
          std::string payload = "{";
          for (auto& kv : assigned) { if (payload.size() > 1) payload += ","; payload += jsonString(kv.first) + ":" + kv.second; }
          annos["assigned-containers"] = payload + "}";
        }
        return;
      }
      if (v_.FitKind == 8) {
        // kunlun/device.go:93-110: the allocated key only, and BAIDU_COM_DEVICE_IDX = the card indices of the LAST
        // container that received devices
        annos[v_.Allocated] = s;
        for (auto& dp : it->second) {
          std::string value;
          for (auto& d : dp) { if (!value.empty()) value += ","; value += std::to_string(d.Idx); }
          if (!value.empty()) annos["BAIDU_COM_DEVICE_IDX"] = value;
        }
        return;
      }
      annos[v_.ToAllocate] = s;
      annos[v_.Allocated] = s;
    }
  }

  bool PolicyNeutralScore() const override { return v_.FitKind == 11; }      // metax/sdevice.go:305-308
  bool PendingQos(const std::string& uuid, std::string& out) const override {
    auto it = sgpu_.jq.find(uuid);
    if (it == sgpu_.jq.end()) return false;
    out = it->second;
    return true;
  }
  void SyncRegistered(const std::vector<DeviceInfo>& devices) const override {      // JitteryQosCache.Sync, metax/qos.go:37-59
    if (v_.FitKind != 11) return;
    for (auto& d : devices) {
      auto it = sgpu_.jq.find(d.ID);
      if (it != sgpu_.jq.end() && it->second == d.QosPolicy) sgpu_.jq.erase(it);
    }
  }

 private:
  Config::PlainVendor v_;
  mutable SgpuState sgpu_;
};

class DCUDevices : public Devices {
 public:
  std::string CommonWord() const override { return "DCU"; }
  FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                const NodeInfo* nodeInfo, const std::string& def, const PodDevices* allocated,
                const QuotaManager* quota) const override {
    FitGates g{Vendor::DCU, "DCU", "hygon.com/use-dcutype", "hygon.com/nouse-dcutype",
               "hygon.com/use-gpuuuid", "hygon.com/nouse-gpuuuid"};
    return fitTemplate(g, devices, request, pod, nodeInfo, def, allocated, quota, GetResourceNames());
  }
  void PatchAnnotations(const Pod&, std::map<std::string, std::string>& annos, const PodDevices& pd) const override {
    // hygon/device.go:241-251
    auto it = pd.find("DCU");
    if (it != pd.end() && !it->second.empty()) {
      std::string s = EncodePodSingleDevice(it->second);
      annos["hami.io/dcu-devices-to-allocate"] = s;
      annos["hami.io/dcu-devices-allocated"] = s;
    }
  }
};

class CambriconDevices : public Devices {
 public:
  std::string CommonWord() const override { return "MLU"; }
  ResourceNames GetResourceNames() const override { return {"MLU/mem", "MLU/core", 256}; }   // cambricon/device.go:60 MemoryFactor
  FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                const NodeInfo* nodeInfo, const std::string& def, const PodDevices* allocated,
                const QuotaManager* quota) const override {
    FitGates g{Vendor::MLU, "MLU", "", "", "cambricon.com/use-gpuuuid", "cambricon.com/nouse-gpuuuid"};
    return fitTemplate(g, devices, request, pod, nodeInfo, def, allocated, quota, GetResourceNames());
  }
  void PatchAnnotations(const Pod&, std::map<std::string, std::string>& annos, const PodDevices& pd) const override {
    // cambricon/device.go:312-332
    auto it = pd.find("MLU");
    if (it != pd.end() && !it->second.empty()) {
      annos["CAMBRICON_DSMLU_ASSIGNED"] = "false";
      for (auto& ctrdevs : it->second) {
        if (!ctrdevs.empty()) {
          const ContainerDevice& d = ctrdevs[0];
          annos["CAMBRICON_DSMLU_PROFILE"] = std::to_string(d.Idx) + "_" + std::to_string(d.Usedcores) + "_" + std::to_string(d.Usedmem / 256);
          break;
        }
      }
      std::string s = EncodePodSingleDevice(it->second);
      annos["hami.io/cambricon-mlu-devices-to-allocate"] = s;
      annos["hami.io/cambricon-mlu-devices-allocated"] = s;
    }
  }
};

class AscendDevices : public Devices {
 public:
  AscendDevices(const AscendConfig& c, bool hamiVnpuCore) : cfg_(c), hamiVnpuCore_(hamiVnpuCore) {
    // InitDevices sorts templates by memory, ascend/device.go:98-100
    std::stable_sort(cfg_.Templates.begin(), cfg_.Templates.end(),
                     [](const AscendTemplate& a, const AscendTemplate& b) { return a.Memory < b.Memory; });
  }
  std::string CommonWord() const override { return cfg_.CommonWord; }
  FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request, const Pod& pod,
                const NodeInfo* nodeInfo, const std::string& def, const PodDevices* allocated,
                const QuotaManager* quota) const override {
    FitGates g{Vendor::ASCEND, cfg_.CommonWord, "", "",
               "hami.io/use-" + cfg_.CommonWord + "-uuid", "hami.io/no-use-" + cfg_.CommonWord + "-uuid", hamiVnpuCore_};
    return fitTemplate(g, devices, request, pod, nodeInfo, def, allocated, quota, GetResourceNames());
  }
  // ascend/device.go:388-424
  float ScoreNode(const PodSingleDevice& podDevices, const std::string&, const NodeInfo*, const std::vector<DeviceUsage>*) const override {
    if (!HasPrefix(cfg_.CommonWord, "Ascend910")) return 0;
    float score = 0;
    for (auto& containerDevices : podDevices) {
      if (containerDevices.empty()) continue;
      std::map<int, int> cntMap;
      for (auto& d : containerDevices) {
        if (!d.HasNetworkID) return 0;
        cntMap[d.NetworkID]++;
      }
      int maxCnt = 0, totalCnt = 0;
      for (auto& kv : cntMap) { if (kv.second > maxCnt) maxCnt = kv.second; totalCnt += kv.second; }
      if (totalCnt == 0) continue;
      score = score + (float)maxCnt / (float)totalCnt;
    }
    return score * (float)10;
  }
  // trimMemory, ascend/device.go:69-79
  std::string trimName(int64_t m) const {
    for (auto& t : cfg_.Templates) if (m <= t.Memory) return t.Name;
    return "";
  }
  void PatchAnnotations(const Pod& pod, std::map<std::string, std::string>& annos, const PodDevices& pd) const override {
    // ascend/device.go:221-253
    auto it = pd.find(cfg_.CommonWord);
    if (it != pd.end() && !it->second.empty()) {
      std::string s = EncodePodSingleDevice(it->second);
      annos["hami.io/" + cfg_.CommonWord + "-devices-to-allocate"] = s;
      annos["hami.io/" + cfg_.CommonWord + "-devices-allocated"] = s;
      annos["predicate-time"] = "";  // canonical rule 6
      std::string vnpuMode;
      if (auto v = anno(pod, "huawei.com/vnpu-mode")) vnpuMode = *v;
      std::string json;
      bool any = false;
      for (auto& dp : it->second) {
        for (auto& val : dp) {
          std::string e = "{";
          bool first = true;
          auto add = [&](const std::string& f) { if (!first) e += ","; e += f; first = false; };
          if (!val.UUID.empty()) add("\"UUID\":\"" + val.UUID + "\"");
          if (vnpuMode != "hami-core") {
            std::string temp = trimName((int64_t)val.Usedmem);
            if (!temp.empty()) add("\"temp\":\"" + temp + "\"");
          }
          if (val.Usedmem != 0) add("\"memory\":" + std::to_string((int64_t)val.Usedmem));
          if (val.Usedcores != 0) add("\"core\":" + std::to_string(val.Usedcores));
          e += "}";
          json += (any ? "," : "") + e;
          any = true;
        }
      }
      annos["huawei.com/" + cfg_.CommonWord] = any ? "[" + json + "]" : "null";
    }
  }

 private:
  AscendConfig cfg_;
  bool hamiVnpuCore_;
};

// getNodeResources, score.go:43-52
std::vector<DeviceUsage*> getNodeResources(NodeUsage& list, const std::string& t) {
  std::vector<DeviceUsage*> l;
  std::string targetType = ToLower(t);
  for (auto& val : list.Devices.DeviceLists)
    if (Contains(ToLower(val.Device.Type), targetType)) l.push_back(&val.Device);
  return l;
}

// isMIGRequest, score.go:54-69
bool isMIGRequest(const ContainerDeviceRequest& request, const std::vector<DeviceUsage*>& devices, const Pod& pod) {
  if (request.Type != "NVIDIA") return false;
  if (auto mode = anno(pod, "nvidia.com/vgpu-mode")) {
    if (!Contains(*mode, "mig")) return false;
  }
  for (auto d : devices) if (d->Mode == "mig") return true;
  return false;
}

}  // namespace

// ---------------------------------------------------------------------------
// Scheduler
// ---------------------------------------------------------------------------
Scheduler::Scheduler(const Config& cfg) : cfg_(cfg) {
  vendors_["NVIDIA"] = std::make_unique<NvidiaGPUDevices>();
  vendors_["DCU"] = std::make_unique<DCUDevices>();
  vendors_["MLU"] = std::make_unique<CambriconDevices>();
  for (auto& a : cfg.Ascend) vendors_[a.CommonWord] = std::make_unique<AscendDevices>(a, cfg.AscendHamiVnpuCore);
  for (auto& p : cfg.Plain) vendors_[p.CommonWord] = std::make_unique<PlainDevices>(p);
}
Scheduler::~Scheduler() = default;

const Devices* Scheduler::Vendor(const std::string& key) const {
  auto it = vendors_.find(key);
  return it == vendors_.end() ? nullptr : it->second.get();
}

void Scheduler::AddNode(const NodeInfo& info) {
  for (auto& kv : info.Devices) { auto v = vendors_.find(kv.first); if (v != vendors_.end()) v->second->SyncRegistered(kv.second); }
  auto it = nodeIndex_.find(info.ID);
  std::vector<DeviceUsage> usage;
  for (auto& kv : info.Devices) {
    for (auto& d : kv.second) {
      DeviceUsage u;  // buildNodeUsage, scheduler.go:665-687
      u.ID = d.ID; u.Index = d.Index; u.Count = d.Count; u.Totalmem = d.Devmem; u.Totalcore = d.Devcore;
      u.Mode = d.Mode; u.Type = d.Type; u.Numa = d.Numa; u.Health = d.Health;
      u.HasNetworkID = d.HasNetworkID; u.NetworkID = d.NetworkID;
      u.MigProfiles = d.MIGProfiles;   // scheduler.go:678
      u.EnfMinor = d.EnfMinor; u.EnfIndex = d.EnfIndex; u.EnfProfiles = d.EnfProfiles; u.EnfMaxSlice = d.EnfMaxSlice;   // CustomInfo, :684
      u.HasSgpu = d.HasSgpu; u.QosPolicy = d.QosPolicy; u.LinkZone = d.LinkZone;
      usage.push_back(u);
    }
  }
  if (usage.size() > 12) throw std::runtime_error("oracle: >12 devices per node (canonical rule 5)");
  if (it == nodeIndex_.end()) {
    auto ne = std::make_unique<NodeEntry>();
    ne->info = info;
    ne->usage = usage;
    nodeIndex_[info.ID] = (int)nodes_.size();
    nodes_.push_back(std::move(ne));
  } else {
    // the reference rebuilds usage on every Filter by joining the pod cache and the node's devices by UUID
    // (getNodesUsage, scheduler.go:757-842): a device that left the annotation and came back finds its pods again
    NodeEntry& ne = *nodes_[it->second];
    ne.info = info;
    ne.usage = usage;
    ne.alive = true;
    for (auto& kv : pods_)
      if (kv.second.NodeID == info.ID) applyUsage(ne, kv.second.Devices, +1, kv.second.AppClass);
    refreshMig(ne);
  }
}

// The MIG half of getNodesUsage (scheduler.go:757-842) for one node, from scratch: which reservations are in use on
// each device, and which devices fail closed (a MIG-mode device holding a pod without a reservation, a reservation
// on a hami-core device, reservations left over for a non-MIG or unknown device).  The result does not depend on
// the order the pods are visited in: reservations are consumed per pod, and Health only ever goes to false.
void Scheduler::refreshMig(NodeEntry& ne) {
  bool any = false;
  for (auto& d : ne.usage) if (d.Mode == "mig") any = true;
  if (!any)
    for (auto& kv : pods_) if (kv.second.NodeID == ne.info.ID && !kv.second.MigAllocs.empty()) { any = true; break; }
  if (!any && !ne.migTouched) return;
  ne.migTouched = any;
  size_t idx = 0;
  for (auto& kv : ne.info.Devices)
    for (auto& d : kv.second) {
      if (idx < ne.usage.size()) { ne.usage[idx].Health = d.Health; ne.usage[idx].MigAllocationsInUse.clear(); }
      idx++;
    }
  for (auto& pkv : pods_) {
    const PodInfo& p = pkv.second;
    if (p.NodeID != ne.info.ID) continue;
    std::map<std::string, std::vector<MigReservation>> byGPU;
    for (auto& a : p.MigAllocs) byGPU[a.GPUUUID].push_back(a);
    for (auto& podsingleds : p.Devices)
      for (auto& ctrdevs : podsingleds.second)
        for (auto& udevice : ctrdevs)
          for (auto& d : ne.usage) {
            if (d.ID != udevice.UUID) continue;
            auto& allocations = byGPU[udevice.UUID];
            if (!allocations.empty()) {
              if (d.Mode == "hami-core") { d.Health = false; continue; }          // scheduler.go:788-792
              d.MigAllocationsInUse.push_back(MigAllocation{allocations[0].Profile, allocations[0].Placement});
              allocations.erase(allocations.begin());
              continue;
            }
            if (d.Mode == "mig") d.Health = false;                                 // :799-802, fail closed
          }
    for (auto& g : byGPU) {                                                        // :810-841
      if (g.second.empty()) continue;
      bool matched = false;
      for (auto& d : ne.usage) {
        if (d.ID != g.first) continue;
        matched = true;
        if (d.Mode != "mig") { d.Health = false; break; }
        for (auto& a : g.second) d.MigAllocationsInUse.push_back(MigAllocation{a.Profile, a.Placement});
        break;
      }
      if (!matched)
        for (auto& d : ne.usage) if (d.Mode == "mig") d.Health = false;
    }
  }
}

bool Scheduler::AddPodWithMig(const std::string& uid, const std::string& ns, const std::string& node, const PodDevices& devices,
                              const std::vector<MigReservation>& allocs) {
  auto it = nodeIndex_.find(node);
  if (it == nodeIndex_.end()) return false;
  auto old = pods_.find(uid);
  std::string appClass;
  if (old != pods_.end()) {
    appClass = old->second.AppClass;
    auto n = nodeIndex_.find(old->second.NodeID);
    if (n != nodeIndex_.end()) {
      applyUsage(*nodes_[n->second], old->second.Devices, -1, old->second.AppClass);
      std::string prevNode = old->second.NodeID;
      pods_.erase(old);
      refreshMig(*nodes_[n->second]);
    } else pods_.erase(old);
  } else {
    quota.AddUsage(ns, CountPodDevices(devices));
  }
  PodInfo pi;
  pi.NodeID = node; pi.Namespace = ns; pi.Devices = devices; pi.MigAllocs = allocs; pi.AppClass = appClass;
  pods_[uid] = pi;
  applyUsage(*nodes_[it->second], devices, +1, appClass);
  refreshMig(*nodes_[it->second]);
  return true;
}

bool Scheduler::RemoveNode(const std::string& name) {
  auto it = nodeIndex_.find(name);
  if (it == nodeIndex_.end() || !nodes_[it->second]->alive) return false;
  nodes_[it->second]->alive = false;
  return true;
}

// The usage a cached pod contributes: scheduler.go:773-785
void Scheduler::applyUsage(NodeEntry& ne, const PodDevices& devs, int sign, const std::string& appClass) {
  for (auto& kv : devs)
    for (auto& ctrdevs : kv.second)
      for (auto& udevice : ctrdevs)
        for (auto& d : ne.usage)
          if (d.ID == udevice.UUID) {
            if (appClass == "online") d.OnlinePods += sign;        // d.Device.PodInfos = append(…, p), scheduler.go:786
            else if (appClass == "offline") d.OfflinePods += sign;
            int32_t slots = std::max(udevice.Slots, 1);
            if (sign > 0) { d.Used = i32add(d.Used, slots); d.Usedmem = i32add(d.Usedmem, udevice.Usedmem); d.Usedcores = i32add(d.Usedcores, udevice.Usedcores); }
            else { d.Used = i32sub(d.Used, slots); d.Usedmem = i32sub(d.Usedmem, udevice.Usedmem); d.Usedcores = i32sub(d.Usedcores, udevice.Usedcores); }
          }
}

bool Scheduler::UsageApply(const std::string& uid, const std::string& node, const std::string& uuid,
                           int32_t slots, int32_t mem, int32_t cores) {
  auto it = nodeIndex_.find(node);
  if (it == nodeIndex_.end()) return false;
  PodInfo& pi = pods_[uid];
  pi.NodeID = node;
  ContainerDevice cd;
  cd.UUID = uuid; cd.Usedmem = mem; cd.Usedcores = cores; cd.Slots = slots;
  PodDevices one;
  one["_"] = PodSingleDevice{ContainerDevices{cd}};
  if (pi.Devices["_"].empty()) pi.Devices["_"].push_back({});
  pi.Devices["_"][0].push_back(cd);
  applyUsage(*nodes_[it->second], one, +1);
  refreshMig(*nodes_[it->second]);
  return true;
}

bool Scheduler::DeletePod(const std::string& uid) {
  auto it = pods_.find(uid);
  if (it == pods_.end()) return false;
  auto n = nodeIndex_.find(it->second.NodeID);
  if (n != nodeIndex_.end()) applyUsage(*nodes_[n->second], it->second.Devices, -1, it->second.AppClass);
  quota.RmUsage(it->second.Namespace, CountPodDevices(it->second.Devices));   // scheduler.go:250-252,1092-1094
  pods_.erase(it);
  if (n != nodeIndex_.end()) refreshMig(*nodes_[n->second]);
  return true;
}

// ---- pkg/device/quota.go ----------------------------------------------------
bool QuotaManager::FitQuota(const std::string& ns, int64_t memreq, const ResourceNames& names, int64_t coresreq) const {
  auto dq = Quotas.find(ns);
  if (dq == Quotas.end()) return true;
  auto m = dq->second.find(names.Memory);
  if (m != dq->second.end()) {
    int64_t limit = m->second.Limit;
    if (names.MemoryFactor > 1) limit = limit * (int64_t)names.MemoryFactor;
    if (m->second.LimitSet && m->second.Used + memreq > limit) return false;
  }
  auto c = dq->second.find(names.Core);
  if (c != dq->second.end() && c->second.LimitSet && c->second.Used + coresreq > c->second.Limit) return false;
  return true;
}
void QuotaManager::AddUsage(const std::string& ns, const std::map<std::string, int64_t>& usage) {
  if (usage.empty()) return;
  auto& dp = Quotas[ns];
  for (auto& kv : usage) dp[kv.first].Used += kv.second;
}
void QuotaManager::RmUsage(const std::string& ns, const std::map<std::string, int64_t>& usage) {
  if (usage.empty()) return;
  auto dp = Quotas.find(ns);
  if (dp == Quotas.end()) return;
  for (auto& kv : usage) {
    auto q = dp->second.find(kv.first);
    if (q == dp->second.end()) continue;
    q->second.Used -= kv.second;
    if (q->second.Used < 0) q->second.Used = 0;
  }
}
void QuotaManager::SetLimit(const std::string& ns, const std::string& name, int64_t limit) {
  Quota& q = Quotas[ns][name];
  q.Limit = limit;
  q.LimitSet = true;
}
void QuotaManager::ClearLimit(const std::string& ns, const std::string& name) {
  auto dq = Quotas.find(ns);
  if (dq == Quotas.end()) return;
  auto q = dq->second.find(name);
  if (q == dq->second.end()) return;
  q->second.Limit = 0;
  q->second.LimitSet = false;
}
bool FitQuotaForDevice(const Scheduler& s, const Pod& pod, const ContainerDevices& tmpDevs, const PodDevices* allocated,
                       const std::string& word, const std::string& devUUID, int64_t memreq, int64_t coresreq) {
  const Devices* v = s.Vendor(word);
  if (!v) return true;    // FitQuota: unknown device name → true (quota.go:63-66)
  return fitQuota(s.quota, v->GetResourceNames(), pod, tmpDevs, allocated, word, devUUID, memreq, coresreq);
}
std::map<std::string, int64_t> Scheduler::CountPodDevices(const PodDevices& pd) const {
  std::map<std::string, int64_t> res;
  for (auto& kv : pd) {
    const Devices* v = Vendor(kv.first);
    if (!v) continue;
    ResourceNames rn = v->GetResourceNames();
    for (auto& ctr : kv.second)
      for (auto& d : ctr) {
        if (!rn.Memory.empty()) res[rn.Memory] += d.Usedmem;
        if (!rn.Core.empty()) res[rn.Core] += d.Usedcores;
      }
  }
  return res;
}

// buildNodeUsage (scheduler.go:654-690) + the usage half of getNodesUsage (:757-842)
NodeUsage Scheduler::BuildNodeUsage(const NodeInfo& info, const Pod& pod) const {
  NodeUsage nu;
  nu.Info = &info;
  nu.Devices.Policy = GetGPUSchedulerPolicyByPod(cfg_.GPUSchedulerPolicy, pod);
  bool nb = false;  // numaBindingRequested, scheduler.go:641-652
  if (auto v = anno(pod, "nvidia.com/numa-bind")) { bool b; if (ParseBool(*v, b) && b) nb = true; }
  nu.Devices.NumaBind = nb;
  auto it = nodeIndex_.find(info.ID);
  const NodeEntry& ne = *nodes_[it->second];
  const std::vector<DeviceUsage>& src = faithful && (size_t)it->second < rebuilt_.size() ? rebuilt_[it->second] : ne.usage;
  for (auto& u : src) {
    DeviceListsScore ds;
    ds.Device = u;
    nu.Devices.DeviceLists.push_back(ds);
  }
  return nu;
}

std::vector<DeviceUsage> Scheduler::Snapshot(const std::string& node) const {
  auto it = nodeIndex_.find(node);
  if (it == nodeIndex_.end()) return {};
  return nodes_[it->second]->usage;
}

// fitInDevices, score.go:81-137
bool Scheduler::FitInDevices(NodeUsage& node, const ContainerDeviceRequests& requests, const Pod& pod,
                             PodDevices& devinput, const DeviceScoringWeights& weights, std::string& reason) const {
  for (auto& dl : node.Devices.DeviceLists) dl.ComputeScore(requests, weights);
  for (auto& kv : requests) {  // canonical rule 4: sorted key order
    const ContainerDeviceRequest& k = kv.second;
    node.Devices.Sort();
    const Devices* devPlugin = Vendor(k.Type);
    if (!devPlugin) { reason = "Device type not found"; return false; }
    std::vector<DeviceUsage*> typeDevices = getNodeResources(node, k.Type);
    if ((int)k.Nums > (int)typeDevices.size() && !isMIGRequest(k, typeDevices, pod)) {
      reason = "NodeInsufficientDevice";
      return false;
    }
    FitResult fr = devPlugin->Fit(typeDevices, k, pod, node.Info, cfg_.GPUSchedulerPolicy, &devinput, &quota);
    if (!fr.fit) { reason = fr.reason; return false; }
    for (auto& val : fr.devs) {
      DeviceUsage* d = nullptr;
      for (auto& v : node.Devices.DeviceLists) if (v.Device.ID == val.UUID) { d = &v.Device; break; }
      if (!d) { reason = "Device with UUID \"" + val.UUID + "\" not found on node " + (node.Info ? node.Info->ID : "") + " after Fit"; return false; }
      if (!devPlugin->AddResourceUsage(*d, val)) {   // score.go:125-128
        reason = "AddResourceUsage failed for device " + d->ID + ": MIG profile and placement allocation failed";
        return false;
      }
    }
    devinput[k.Type].push_back(fr.devs);
  }
  return true;
}

// scoreNode, score.go:294-346 (+ allocateInitContainers :226-254, allocateAppContainers :256-286,
// snapshot/update/applyPeakUsage :154-224; the peak map only feeds the discarded original node)
NodeScoreResult Scheduler::ScoreNode(const std::string& nodeID, NodeUsage& node, const Pod& pod,
                                     const std::string& userNodePolicy, const DeviceScoringWeights& weights) const {
  NodeScoreResult res;
  std::set<std::string> baseTypes;  // nodeDeviceBaseTypes, score.go:71-79
  for (auto& dl : node.Devices.DeviceLists) baseTypes.insert(dl.Device.Type);
  int numInit = pod.NumInitContainers;

  PodDevices initAllocs;
  bool haveInit = false;
  if (numInit > 0) {
    haveInit = true;
    for (int i = 0; i < (int)pod.Reqs.size(); i++) {
      if (i >= numInit) break;
      const ContainerDeviceRequests& req = pod.Reqs[i];
      if (req.empty()) {
        for (auto& typ : baseTypes) initAllocs[typ].push_back({});
        continue;
      }
      NodeUsage nodeCopy = node;
      std::string reason;
      if (!FitInDevices(nodeCopy, req, pod, initAllocs, weights, reason)) { res.fit = false; res.reason = reason; return res; }
      for (auto& typ : baseTypes)
        if ((int)initAllocs[typ].size() == i) initAllocs[typ].push_back({});
    }
  }

  NodeUsage appNodeCopy = node;
  NodeScore score;
  score.NodeID = nodeID;
  score.Score = ComputeDefaultScore(appNodeCopy.Devices);
  std::vector<DeviceUsage> snapshot;        // score.SnapshotDevice, score.go:328
  for (auto& dl : appNodeCopy.Devices.DeviceLists) snapshot.push_back(dl.Device);

  int appIndex = 0;
  for (int ctrid = 0; ctrid < (int)pod.Reqs.size(); ctrid++) {
    if (ctrid < numInit) continue;
    const ContainerDeviceRequests& n = pod.Reqs[ctrid];
    int sums = 0;
    for (auto& kv : n) sums += (int)kv.second.Nums;
    if (sums == 0) {
      for (auto& typ : baseTypes) score.Devices[typ].push_back({});
      appIndex++;
      continue;
    }
    std::string reason;
    if (!FitInDevices(appNodeCopy, n, pod, score.Devices, weights, reason)) { res.fit = false; res.reason = reason; return res; }
    for (auto& typ : baseTypes)
      if ((int)score.Devices[typ].size() == appIndex) score.Devices[typ].push_back({});
    appIndex++;
  }

  if (numInit > 0 && haveInit) {
    for (auto& kv : initAllocs) {
      PodSingleDevice merged = kv.second;
      auto& cur = score.Devices[kv.first];
      merged.insert(merged.end(), cur.begin(), cur.end());
      cur = merged;
    }
  }

  // OverrideScore, policy/node_policy.go:67-95 (sorted key order)
  float devScore = 0;
  for (auto& kv : score.Devices) {
    const Devices* dev = Vendor(kv.first);
    if (!dev) continue;
    float sc = dev->ScoreNode(kv.second, userNodePolicy, node.Info, &snapshot);
    if (dev->PolicyNeutralScore()) sc = (userNodePolicy == "spread" ? (float)-10000 : (float)10000) * sc;     // node_policy.go:82-88
    devScore = devScore + sc;
  }
  score.Score = score.Score + devScore;
  res.fit = true;
  res.score = score;
  return res;
}

void Scheduler::scoreRange(const std::vector<int>& cand, size_t lo, size_t hi, const Pod& pod,
                           const std::string& userNodePolicy, const DeviceScoringWeights& w,
                           std::vector<NodeScoreResult>& out) const {
  for (size_t i = lo; i < hi; i++) {
    if (cand[i] < 0) continue;
    const NodeEntry& ne = *nodes_[cand[i]];
    NodeUsage nu = BuildNodeUsage(ne.info, pod);
    out[i] = ScoreNode(ne.info.ID, nu, pod, userNodePolicy, w);
  }
}

// Scheduler.Filter, scheduler.go:1067-1154
FilterResult Scheduler::Filter(const Pod& pod, const std::vector<std::string>* nodeNames) {
  FilterResult fr;
  bool hasHAMiResource = false;
  for (auto& reqMap : pod.Reqs) if (!reqMap.empty()) { hasHAMiResource = true; break; }
  if (!hasHAMiResource) { fr.status = 2; return fr; }

  // TakeAndDeletePod, scheduler.go:1092-1094
  DeletePod(pod.UID);

  std::vector<int> cand;
  std::vector<std::string> names;
  if (nodeNames) {
    for (auto& n : *nodeNames) {
      auto it = nodeIndex_.find(n);
      names.push_back(n);
      cand.push_back(it == nodeIndex_.end() || !nodes_[it->second]->alive ? -1 : it->second);
    }
  } else {
    for (size_t i = 0; i < nodes_.size(); i++)
      if (nodes_[i]->alive) { cand.push_back((int)i); names.push_back(nodes_[i]->info.ID); }
  }

  // calcScoreWithOptions, score.go:360-419
  DeviceScoringWeights weights;
  if (auto v = anno(pod, "hami.io/device-scoring-weights")) {
    std::string err;
    if (!ParseDeviceScoringWeights(*v, weights, err)) { fr.status = -4; fr.error = err; return fr; }
  }
  std::string userNodePolicy = cfg_.NodeSchedulerPolicy;  // resolveNodeSchedulerPolicy, score.go:139-146
  if (auto v = anno(pod, "hami.io/node-scheduler-policy")) userNodePolicy = *v;

  if (faithful) {
    // getNodesUsage, scheduler.go:744-842: fresh usage for EVERY registered node, then replay every cached pod
    rebuilt_.assign(nodes_.size(), {});
    for (size_t i = 0; i < nodes_.size(); i++) {
      rebuilt_[i] = nodes_[i]->usage;  // deep copy per node (nodes.go:150-170)
      for (auto& d : rebuilt_[i]) { d.Used = 0; d.Usedmem = 0; d.Usedcores = 0; }
    }
    for (auto& kv : pods_) {
      auto n = nodeIndex_.find(kv.second.NodeID);
      if (n == nodeIndex_.end()) continue;
      for (auto& pd : kv.second.Devices)
        for (auto& ctrdevs : pd.second)
          for (auto& udevice : ctrdevs)
            for (auto& d : rebuilt_[n->second])
              if (d.ID == udevice.UUID) {
                int32_t slots = std::max(udevice.Slots, 1);
                d.Used = i32add(d.Used, slots);
                d.Usedmem = i32add(d.Usedmem, udevice.Usedmem);
                d.Usedcores = i32add(d.Usedcores, udevice.Usedcores);
              }
    }
  }
  std::vector<NodeScoreResult> results(cand.size());
  int T = std::max(1, std::min(threads, (int)cand.size()));
  if (T == 1) {
    scoreRange(cand, 0, cand.size(), pod, userNodePolicy, weights, results);
  } else {
    std::vector<std::thread> th;
    size_t chunk = (cand.size() + T - 1) / T;
    for (int t = 0; t < T; t++) {
      size_t lo = t * chunk, hi = std::min(cand.size(), lo + chunk);
      if (lo >= hi) break;
      th.emplace_back([&, lo, hi] { scoreRange(cand, lo, hi, pod, userNodePolicy, weights, results); });
    }
    for (auto& t : th) t.join();
  }

  // sort.Sort(nodeScores) + take last (scheduler.go:1117-1118) under canonical rules 2-3
  int best = -1;
  bool spread = userNodePolicy == "spread";
  for (size_t i = 0; i < cand.size(); i++) {
    if (cand[i] < 0) { fr.failedNodes.push_back({names[i], "node unregistered"}); continue; }
    if (!results[i].fit) { fr.failedNodes.push_back({names[i], results[i].reason}); continue; }
    if (best < 0) { best = (int)i; continue; }
    float s = results[i].score.Score, b = results[best].score.Score;
    if (spread ? (s <= b) : (s >= b)) best = (int)i;
  }
  if (best < 0) { fr.status = 1; return fr; }

  fr.status = 0;
  fr.nodeIndex = best;
  fr.chosen = results[best].score;
  const NodeScore& m = fr.chosen;
  fr.annotations["hami.io/vgpu-node"] = m.NodeID;
  fr.annotations["hami.io/vgpu-time"] = "";  // canonical rule 6
  for (auto& kv : vendors_) kv.second->PatchAnnotations(pod, fr.annotations, m.Devices);

  // commit: scheduler.go:1132-1138
  PodDevices effective = CollapseInitContainerUsage(pod.NumInitContainers, m.Devices);
  PodInfo pi;
  pi.NodeID = m.NodeID;
  pi.Namespace = pod.Namespace;
  pi.Devices = effective;
  quota.AddUsage(pod.Namespace, CountPodDevices(effective));   // scheduler.go:1136-1138
  if (auto cls = anno(pod, "metax-tech.com/sgpu-app-class")) pi.AppClass = *cls;
  applyUsage(*nodes_[cand[best]], effective, +1, pi.AppClass);
  pods_[pod.UID] = pi;      // args.Pod carries no hami.io/vgpu-mig-allocations yet: a MIG pick fails closed until the informer delivers it
  refreshMig(*nodes_[cand[best]]);
  return fr;
}

}  // namespace oracle
