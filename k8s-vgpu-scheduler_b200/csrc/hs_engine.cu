// hs_engine.cu — host side of the B200 Filter/Score engine and its C ABI
// (include/hami_sched.h).  It is the C++ mirror of the pieces of pkg/scheduler the
// hot path touches: nodeManager (inventory), podManager (what each pod holds),
// the request/annotation plumbing around Filter, and the result encoding.  Strings
// (types, UUIDs, policies, annotations) are resolved to integers HERE; the kernels in
// hs_kernels.cu only see the SoA inventory and integer request descriptors.
// There is no CPU fallback: without a CUDA device hs_create fails.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <deque>
#include <future>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <set>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../../include/hami_sched.h"
#include "hs_ingest.h"
#include "hs_types.cuh"

extern "C" int hs_launch_decide(const KernelArgs* ka, int grid, int block, cudaStream_t stream, int fast_only);
extern "C" int hs_decide_max_blocks(int block, int device, int fast_only);
extern "C" int hs_launch_sweep(const KernelArgs* ka, int op, int mode, int grid, int block, cudaStream_t stream);
extern "C" int hs_sweep_blocks_per_sm(int mode);
extern "C" int hs_launch_shard_decide(const KernelArgs* ka, int op, cudaStream_t stream);
extern "C" int hs_launch_shard_explain(const KernelArgs* ka, int op, int grid, cudaStream_t stream);

namespace {

struct BatchJob;
std::string g_create_error;

// ---- Go string helpers (ASCII subset, as the reference's inputs are) ------------------------
bool isSpace(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }
std::string trimSpace(const std::string& s) {
  size_t a = 0, b = s.size();
  while (a < b && isSpace((unsigned char)s[a])) a++;
  while (b > a && isSpace((unsigned char)s[b - 1])) b--;
  return s.substr(a, b - a);
}
std::vector<std::string> splitComma(const std::string& s) {
  std::vector<std::string> out;
  size_t start = 0;
  for (;;) {
    size_t p = s.find(',', start);
    if (p == std::string::npos) { out.push_back(s.substr(start)); break; }
    out.push_back(s.substr(start, p - start));
    start = p + 1;
  }
  return out;
}
std::string upper(std::string s) { for (auto& c : s) if (c >= 'a' && c <= 'z') c = (char)(c - 32); return s; }
std::string lower(std::string s) { for (auto& c : s) if (c >= 'A' && c <= 'Z') c = (char)(c + 32); return s; }
bool parseBoolTrue(const std::string& s) {  // strconv.ParseBool(s) == (true, nil)
  return s == "1" || s == "t" || s == "T" || s == "TRUE" || s == "true" || s == "True";
}
bool parseBool(const std::string& s, bool& v) {  // strconv.ParseBool
  if (parseBoolTrue(s)) { v = true; return true; }
  if (s == "0" || s == "f" || s == "F" || s == "FALSE" || s == "false" || s == "False") { v = false; return true; }
  return false;
}
bool policyContains(const std::string& policy, const char* name) {  // util/util.go:290-298
  for (auto& p : splitComma(policy)) if (trimSpace(p) == name) return true;
  return false;
}

// device.CheckType, devices.go:719-738, for one card type
bool checkType(const std::string* use, const std::string* nouse, const std::string& cardTypeUpper) {
  auto match = [&](const std::string& list) {
    for (auto& t0 : splitComma(list)) {
      std::string t = trimSpace(t0);
      if (!t.empty() && cardTypeUpper.find(upper(t)) != std::string::npos) return true;
    }
    return false;
  };
  if (use && !trimSpace(*use).empty() && !match(*use)) return false;
  if (nouse && !trimSpace(*nouse).empty() && match(*nouse)) return false;
  return true;
}

// util.ParseDeviceScoringWeights, util/device_scoring_weights.go:58-102
// `why` receives the text after "invalid hami.io/device-scoring-weights annotation <value>: " (fmt's %q on the key where the
// reference names it)
std::string quoteGo(const std::string& s) {       // strconv.Quote for what an annotation can hold
  std::string out = "\"";
  for (unsigned char c : s) {
    if (c == '"') out += "\\\"";
    else if (c == '\\') out += "\\\\";
    else if (c == '\n') out += "\\n";
    else if (c == '\t') out += "\\t";
    else if (c == '\r') out += "\\r";
    else if (c < 0x20 || c == 0x7f) { char b[8]; snprintf(b, sizeof b, "\\x%02x", c); out += b; }
    else out += (char)c;
  }
  return out + "\"";
}
bool parseWeights(const std::string& value, int64_t w[3], std::string& why) {
  w[0] = w[1] = w[2] = 0;
  auto parts = splitComma(value);
  if (parts.size() != 3) { why = "expected slot, core, and memory weights"; return false; }
  std::set<std::string> seen;
  for (auto& part : parts) {
    std::string p = trimSpace(part);
    size_t eq = p.find('=');
    if (eq == std::string::npos) { why = "expected key=value entries"; return false; }
    std::string key = trimSpace(p.substr(0, eq)), val = trimSpace(p.substr(eq + 1));
    if (seen.count(key)) { why = "duplicate " + quoteGo(key) + " weight"; return false; }
    const std::string notInt = quoteGo(key) + " weight must be an integer";
    if (val.empty()) { why = notInt; return false; }
    size_t i = 0;
    bool neg = false;
    if (val[0] == '+' || val[0] == '-') { neg = val[0] == '-'; i = 1; }
    if (i >= val.size()) { why = notInt; return false; }
    unsigned long long acc = 0;
    for (; i < val.size(); i++) {
      if (val[i] < '0' || val[i] > '9') { why = notInt; return false; }
      unsigned long long d = (unsigned long long)(val[i] - '0');
      if (acc > (0x8000000000000000ULL - d) / 10) { why = notInt; return false; }
      acc = acc * 10 + d;
    }
    if (!neg && acc > 0x7fffffffffffffffULL) { why = notInt; return false; }
    if (neg && acc != 0) { why = quoteGo(key) + " weight must not be negative"; return false; }
    int64_t v = (int64_t)acc;
    if (key == "slot") w[0] = v; else if (key == "core") w[1] = v; else if (key == "memory") w[2] = v;
    else { why = "unknown weight " + quoteGo(key); return false; }
    seen.insert(key);
  }
  if (w[0] == 0 && w[1] == 0 && w[2] == 0) { why = "at least one weight must be positive"; return false; }
  return true;
}

struct DevSlot {
  std::string uuid, type, mode, vendor;
  uint32_t index = 0;
  int32_t count = 0, devmem = 0, devcore = 0, numa = 0, network_id = -1;
  bool health = true;
  int type_id = 0, mode_id = 0, vendor_id = 0;
  bool sgpu = false;        // a metax sGPU device (DF_SGPU)
  int mig_geom = 0;         // index into hs_handle::mig_geoms for a GPU in MIG mode (0 = the empty geometry); enflame: the node's profile table
  std::string enf_minor, enf_index;                               // enflame CustomInfo["minor"] / ["index"]
  std::vector<std::pair<std::string, std::string>> enf_profiles;  // … ["profiles"]
  int enf_max_slice = 0;                                          // … ["maxSlice"]
};

// One distinct MigProfiles list (device.MigProfile, devices.go:79-90) and what the device needs of it (MigGeom)
struct MigGeomHost {
  struct Prof { std::string name; int32_t memory_mb, core; uint32_t slice_count; std::vector<std::pair<uint32_t, uint32_t>> placements; };
  std::vector<Prof> profiles;   // as registered
  // enflame: the node's DRS profile candidates in selectProfileByRequest's order (enflame/device.go:567-619)
  struct Enf { std::string name, id; long long size, mem_gb, core; };
  std::vector<Enf> enf;
  bool enflame = false;
  MigGeom dev;
};
struct MigRes { std::string uuid, profile; uint32_t start = 0, size = 0; };   // one entry of hami.io/vgpu-mig-allocations

struct NodeSlot {
  std::string name;
  std::vector<DevSlot> devs;
  bool alive = false;
  bool has_cordon = false;
  std::string cordon;
  int vnpu = -1;
  std::vector<std::string> base_types;  // nodeDeviceBaseTypes, score.go:71-79 (sorted, unique)
  std::vector<std::string> vendor_base_types;   // the base types that are themselves a registered vendor word: their
                                                // PatchAnnotations sees a (padded) entry under its key even when the pod asked for none
  bool has_mig = false;        // some device is in MIG mode
  bool mig_touched = false;    // the last mig_refresh left MIG bookkeeping in the node's dflags
};

struct VendorInfo {
  int id;
  std::string word;
  std::string use_type, nouse_type, use_uuid, nouse_uuid;
  std::string to_allocate, allocated;   // InRequestDevices / SupportDevices annotation keys
  bool ascend = false, ascend910 = false;
  int gate = 0;             // device-side gate class (HV_*): plain vendors share hygon's gate list (HV_DCU)
  bool count_only = false;  // biren's Fit: no memory / core gates (EF_COUNT_ONLY)
  bool whole_card = false;  // enflame GCU's Fit: a card in use at all is a conflict (EF_WHOLE_CARD)
  bool mthreads = false;    // mthreads' Fit: template gates + customFilterRule; its own annotation set (EF_MTHREADS)
  bool neuron = false;      // AWS Neuron: core bit masks, graphSelect for several devices, per-pod awsusageinfo state (EF_NEURON)
  bool metax = false;       // metax whole GPUs: mthreads' gates without uuid lists + ScoreNode from the node's annotations (EF_METAX)
  int fit_kind = 0;         // HS_FIT_* as registered (hs_filter_simulate hands it on to the child handle)
  bool vastai = false;      // vastai: count-only gates + die mode (EF_VASTAI)
  bool sgpu = false;        // metax sGPU: application classes, QoS switches, MetaLink prioritisation, policy-neutral ScoreNode (EF_SGPU)
  int option = 0;           // hs_vendor_config.option (sGPU: MetaxConfig.TopologyAware)
  bool enflame = false;     // enflame shared vGCU: DRS profile selection per request (EF_ENFLAME)
  bool kunlun = false;      // kunlun whole XPUs: graghSelect, slot-only usage, wing ScoreNode (EF_KUNLUN)
  bool vxpu = false;        // … the shared "XPU" type on the same graghSelect (EF_KUNLUN | EF_VXPU)
  bool amd = false;         // amd's Fit: whole-card defaults, per-device core share, picks carry the device's type (EF_AMD)
  std::vector<std::pair<std::string, int64_t>> templates;  // sorted by memory
};

struct PodRec {           // podManager entry: what to subtract when the pod goes away
  int node;
  int4 delta[HS_D];
  bool init_released = false;   // PodInfo.InitContainerResourceReleased (pods.go:40-53)
  std::string ns;               // pod.Namespace, for quotaManager.RmUsage
  std::vector<std::pair<std::string, int4>> parked;   // usage on devices that left the node's register annotation, by UUID
  uint8_t n_ent[HS_D] = {0, 0, 0, 0, 0, 0, 0, 0};     // cached ContainerDevice entries per device slot (getNodesUsage visits each one)
  std::vector<MigRes> mig;                            // decoded hami.io/vgpu-mig-allocations, annotation order
  uint8_t app_class = 0;                              // metax-tech.com/sgpu-app-class of the pod: 1 online, 2 offline (DeviceUsage.PodInfos)
};

struct EntryMeta {
  std::string type;       // k.Type == PodDevices key
  int container;
  int vendor;
};

struct OpMeta {           // host-side memory of one op of the current batch
  int kind = OP_NOP;      // OP_FILTER / OP_RELEASE / OP_NOP
  int status = HS_OK;     // compile-time verdict: HS_PASSTHROUGH / HS_ERR_*
  std::string uid, vnpu_mode, ns;
  int sg_class = 0, sg_qos = 0;         // metax sGPU: the pod's application class (0 none, 1 online, 2 offline, 3 other) and QoS id
  std::vector<std::string> ctr_names;   // Container.Name by position (enflame's assigned-containers annotation)
  int n_ctr = 0, n_init = 0;
  std::vector<EntryMeta> entries;
  std::vector<uint8_t> ctr_begin;
  int user_op = -1;       // index of the caller's op this device op belongs to
  bool dry = false;       // filterSimulation: nothing of this op is kept
};

template <class T>
struct DBuf {             // device buffer + pinned host mirror
  T* d = nullptr;
  T* h = nullptr;
  size_t cap = 0;
  bool ensure(size_t n) {
    if (n <= cap) return true;
    size_t ncap = std::max<size_t>(n, cap * 2);
    if (ncap < 16) ncap = 16;
    T* nd = nullptr; T* nh = nullptr;
    if (cudaMalloc(&nd, ncap * sizeof(T)) != cudaSuccess) return false;
    if (cudaMallocHost(&nh, ncap * sizeof(T)) != cudaSuccess) { cudaFree(nd); return false; }
    if (d) cudaFree(d);
    if (h) cudaFreeHost(h);
    d = nd; h = nh; cap = ncap;
    return true;
  }
  void release() { if (d) cudaFree(d); if (h) cudaFreeHost(h); d = nullptr; h = nullptr; cap = 0; }
};

}  // namespace

struct hs_handle {
  std::string err;
  std::string node_policy = "binpack", gpu_policy = "spread";
  bool ascend_vnpu_default = false;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // chained launches of a pipelined batch: results come back on a second stream while the next launch already runs
  cudaStream_t stream2 = nullptr, stream3 = nullptr;      // results back / request records up
  cudaEvent_t cu[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ck0[4] = {nullptr, nullptr, nullptr, nullptr}, ck1[4] = {nullptr, nullptr, nullptr, nullptr}, cc[4] = {nullptr, nullptr, nullptr, nullptr};
  uint32_t* d_chain = nullptr;          // 2 x {ops completed, fail cursor}, alternating between launches (KernelArgs::chain_in/out)
  uint32_t* h_chain = nullptr;          // pinned staging for resets
  double last_ms = 0;
  uint64_t launches = 0;
  int block = 256;          // threads per CTA of the decide kernel: 8 lanes per node
  bool small_direct = false;    // this batch: results are written straight into the pinned host mirrors (device_begin)
  bool no_small_direct = false; // HS_NO_SMALL_DIRECT=1: always copy results back (A/B of the single-call latency)
  bool no_fast = false;     // HS_NO_FAST=1: force the general evaluator (A/B testing of the two device paths)
  bool no_resident = false; // HS_NO_RESIDENT=1: force the streaming mode of the decide loop (tests: the lane-dealt general path on small inventories)

  std::vector<VendorInfo> vendors;
  std::unordered_map<std::string, int> vendor_by_word;
  std::vector<std::string> types, types_lower, types_upper;
  std::unordered_map<std::string, int> type_ids;
  std::vector<std::string> modes;
  // dynamic MIG: distinct profile lists, their device table, and whether any GPU in MIG mode is registered
  std::deque<MigGeomHost> mig_geoms;
  MigGeom* d_mig_tab = nullptr;
  size_t mig_tab_uploaded = 0;
  bool has_mig = false;
  size_t n_mig_pods = 0;       // cached pods that carry reservations

  std::vector<NodeSlot> nodes;
  std::deque<std::string> name_store;      // node names at stable addresses (slot order)
  std::vector<const char*> name_ptr;       // name_store[slot].c_str()
  std::unordered_map<std::string, int> node_by_name;
  std::unordered_map<std::string, std::vector<uint32_t>> uuid_index;  // uuid -> (slot<<8 | dev)
  int n_alive = 0;
  bool has_dead = false;

  // inventory mirrors (host) + device arrays
  std::vector<int4> h_dyn, h_stat;
  std::vector<uint32_t> h_flags;
  std::vector<int4> h_tot;          // per node {flags, sum count, sum totalcore, sum totalmem}: the streamed sweep's node record
  int4* d_tot = nullptr;
  std::vector<int16_t> h_pair;
  int4* d_dyn = nullptr; int4* d_stat = nullptr; uint32_t* d_flags = nullptr; int16_t* d_pair = nullptr;
  int cap_nodes = 0;       // slots allocated (multiple of HS_TILE)
  int dirty_lo = 1 << 30, dirty_hi = -1;   // tile range to upload before the next launch

  std::unordered_map<std::string, PodRec> pods;

  // batch staging
  DBuf<DevPod> b_pods;
  DBuf<DevEntry> b_entries;
  DBuf<uint32_t> b_uuid;
  DBuf<int4> b_rel;
  DBuf<DevDecision> b_dec;
  DBuf<DevAlloc> b_alloc;
  DBuf<int4> b_delta;
  DBuf<OpSync> b_best;
  DBuf<DevFail> b_fail;
  DBuf<uint32_t> b_counters;
  DBuf<int32_t> b_candpos;
  // windows (hs_window.cuh): per-tile sorted keys / per-op fit masks / heads, sized with the inventory
  unsigned long long* w_skey = nullptr; uint32_t* w_rmask = nullptr; unsigned long long* w_head = nullptr; uint8_t* w_rank = nullptr;
  uint32_t* w_stats = nullptr;
  int w_cap_nodes = 0;
  bool windows = false;                // HS_WINDOW=1: decide runs of common-shape ops in windows (hs_window.cuh); measured slower than the one-op path on C3, so off by default
  DBuf<unsigned long long> b_timing;   // HS_TIMING=1: phase clocks of block 0
  DBuf<unsigned long long> b_trace;    // HS_TRACE=1: per-op arrival times of every CTA for the first ops of each launch
  bool trace = false;
  static constexpr int kTraceOps = 512, kTraceStride = 320;
  bool timing = false;
  // sharded mode (one process per GPU): contiguous node block per rank, NCCL allreduce-max of the packed key per decision
  int shard_rank = 0, shard_world = 1;
  int64_t pos_base = 0;
  void* nccl_comm = nullptr;
  // fused path: peer-mapped mailboxes (cudaIpc) for the in-kernel arg-best exchange over NVLink
  unsigned long long* d_mbox = nullptr;       // this rank's mailbox [4 rows][8 ranks] + xresult[4] + xerror
  unsigned long long* mbox_peer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  bool xgpu_connected = false;
  bool shard_use_nccl = false;               // HS_SHARD_NCCL=1: keep the launch-per-decision NCCL baseline
  uint32_t epoch_base = 0;
  bool force_shard = false;  // HS_FORCE_SHARD=1: run the sharded pipeline (sweep / decide / explain kernels) with world == 1
  struct FailTab { uint32_t k0, k1, k2, k3; const std::string* s; };
  static constexpr uint32_t kFailTab = 8192;
  std::vector<FailTab> fail_tab = std::vector<FailTab>(kFailTab, FailTab{0, 0, 0, 0, nullptr});
  // device.QuotaManager (quota.go): (namespace, vendor id) -> {memory, cores}; Used is tracked for every namespace
  // (addUsageLocked auto-creates entries), rows with a limit set travel to the device with each batch
  struct QuotaEnt { int64_t limit[2] = {0, 0}; bool set[2] = {false, false}; int64_t used[2] = {0, 0}; };
  struct NsQuota { QuotaEnt v[32]; uint32_t present = 0; };      // per namespace: one entry per vendor id, `present` = entry exists
  std::unordered_map<std::string, NsQuota> quota;
  QuotaEnt* quota_find(const std::string& ns, int vid) {
    auto it = quota.find(ns);
    return (it == quota.end() || !((it->second.present >> vid) & 1u)) ? nullptr : &it->second.v[vid];
  }
  QuotaEnt& quota_at(const std::string& ns, int vid) {
    NsQuota& q = quota[ns];
    q.present |= 1u << vid;
    return q.v[vid];
  }
  int32_t mem_factor[32] = {1, 1, 256, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};   // ResourceNames.MemoryFactor; MLU = 256 (cambricon/device.go:60)
  bool any_quota_limit = false;
  NsQuota* last_quota = nullptr;
  std::string last_quota_ns;
  std::vector<std::pair<std::string, int>> quota_rows;           // rows of the current batch's device table
  DBuf<long long> b_quota;
  hs_handle* sim = nullptr;             // child handle holding the transient inventory of hs_filter_simulate
  bool dry_run = false;                 // set around hs_filter_dry: ops neither release nor commit
  uint64_t bytes_h2d = 0, bytes_d2h = 0;  // measured host<->device copy volume of the batch path
  std::unique_ptr<BatchJob> job;  // the batch of the last hs_run_batch (compile state + pipeline)
  size_t n_parked = 0;                  // PodRec::parked entries alive in h->pods
  bool poisoned = false;                // a CUDA call failed while a batch was on the device: HBM and the host mirror may differ
  bool hostprof = false;                // HS_HOSTPROF=1: split the summary time by result kind, printed at hs_destroy
  double prof_s[4] = {0, 0, 0, 0};      // fit / no-fit / other / count
  double host_s[4] = {0, 0, 0, 0};      // diagnostic: cumulative host seconds in compile / device+copies / absorb / summary
  uint32_t fail_cap = 32;
  uint32_t pipe_chunk = 512;
  int sync2_base = 0;                   // first quota-barrier slot in b_best (== op capacity of the current batch)
  int grid_now = 1;                     // grid of the decide kernel for the current batch
  int maxb_cache[2] = {0, 0}, maxb_block = 0;   // co-resident CTAs of the two decide kernels for `maxb_block` threads (queried once)
  int grid_fast = 0;                    // … of its instantiation without the general evaluator (0 = not used: HS_NO_FAST_KERNEL, block != 256)
  bool fast_kernel = true;
  uint32_t fail_cap_now = 32;           // no-fit decisions one launch can hold before the host drains the reasons
  bool fail_cap_fixed = false;          // HS_FAIL_CAP given: use it as is (tests of the drain-and-resume path)
  std::vector<OpMeta> metas;            // per device op
  std::vector<int> user_first_op;       // per caller op: index of its FILTER/RELEASE device op
  std::vector<int> user_status;
  std::vector<std::string> cand_names;  // explicit candidate list of the last hs_filter
  std::vector<int> cand_slot;
  bool have_cand = false;
  std::deque<std::vector<DevFail>> fail_segs;   // drained fail records, one segment per drain: [slot][n_nodes] (segments never move)
  std::vector<const DevFail*> op_fail_row;      // per device op: its [n_nodes] reason records, or NULL

  // mask cache
  std::unordered_map<std::string, std::array<uint64_t, 4>> mask_cache;
  size_t mask_cache_types = 0;

  // result arena (valid until the next call)
  std::deque<std::string> r_strs;
  std::vector<hs_alloc_entry> r_entries;
  std::vector<hs_kv> r_annos;
  std::vector<hs_failed_node> r_failed;
  std::string r_buf;                    // device-list annotation of the last assembled result
  DevFail last_fail = {0xffffffffu, 0, 0, 0};
  const std::string* last_fail_str = nullptr;
  std::map<std::array<uint32_t, 4>, std::string> reason_cache;

  const char* keep(const std::string& s) { r_strs.push_back(s); return r_strs.back().c_str(); }
};

namespace {

#define CUDA_OK(h, call)                                                                      \
  do {                                                                                        \
    cudaError_t _e = (call);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      (h)->err = std::string(#call) + ": " + cudaGetErrorString(_e);                          \
      return HS_ERR_CUDA;                                                                     \
    }                                                                                         \
  } while (0)

static const char* kReasonNames[HS_R_COUNT] = {
    "CardTypeMismatch", "CardUuidMismatch", "CardTimeSlicingExhausted", "CardComputeUnitsExhausted",
    "CardInsufficientMemory", "CardInsufficientCore", "CardNotHealth", "CardCordoned", "NumaNotFit",
    "ExclusiveDeviceAllocateConflict", "CardNotFoundCustomFilterRule", "CardMigTopologyInfeasible",
    "AllocatedCardsInsufficientRequest", "ResourceQuotaNotFit", "ModeNotFit",
    "CardUnhealthy", "CardAppClassMismatch", "CardQosPolicyMismatch"};

void add_vendor(hs_handle* h, const std::string& word, bool ascend) {
  VendorInfo v;
  v.id = (int)h->vendors.size();
  v.word = word;
  v.ascend = ascend;
  if (word == "NVIDIA") {
    v.use_type = "nvidia.com/use-gputype"; v.nouse_type = "nvidia.com/nouse-gputype";
    v.use_uuid = "nvidia.com/use-gpuuuid"; v.nouse_uuid = "nvidia.com/nouse-gpuuuid";
    v.to_allocate = "hami.io/vgpu-devices-to-allocate"; v.allocated = "hami.io/vgpu-devices-allocated";
  } else if (word == "DCU") {
    v.use_type = "hygon.com/use-dcutype"; v.nouse_type = "hygon.com/nouse-dcutype";
    v.use_uuid = "hygon.com/use-gpuuuid"; v.nouse_uuid = "hygon.com/nouse-gpuuuid";
    v.to_allocate = "hami.io/dcu-devices-to-allocate"; v.allocated = "hami.io/dcu-devices-allocated";
  } else if (word == "MLU") {
    v.use_uuid = "cambricon.com/use-gpuuuid"; v.nouse_uuid = "cambricon.com/nouse-gpuuuid";
    v.to_allocate = "hami.io/cambricon-mlu-devices-to-allocate"; v.allocated = "hami.io/cambricon-mlu-devices-allocated";
  } else {
    v.use_uuid = "hami.io/use-" + word + "-uuid"; v.nouse_uuid = "hami.io/no-use-" + word + "-uuid";
    v.to_allocate = "hami.io/" + word + "-devices-to-allocate"; v.allocated = "hami.io/" + word + "-devices-allocated";
    v.ascend910 = word.compare(0, 9, "Ascend910") == 0;
  }
  v.gate = v.id;
  h->vendor_by_word[word] = v.id;
  h->vendors.push_back(v);
}

int intern_type(hs_handle* h, const std::string& t) {
  auto it = h->type_ids.find(t);
  if (it != h->type_ids.end()) return it->second;
  if (h->types.size() >= HS_MAX_TYPES) return -1;
  int id = (int)h->types.size();
  h->types.push_back(t);
  h->types_lower.push_back(lower(t));
  h->types_upper.push_back(upper(t));
  h->type_ids[t] = id;
  return id;
}

int intern_mode(hs_handle* h, const std::string& m) {
  for (size_t i = 0; i < h->modes.size(); i++) if (h->modes[i] == m) return (int)i;
  if (h->modes.size() >= 8) return -1;
  h->modes.push_back(m);
  return (int)h->modes.size() - 1;
}

int grow_inventory(hs_handle* h, int need_nodes) {
  if (need_nodes <= h->cap_nodes) return HS_OK;
  int ncap = std::max(need_nodes, h->cap_nodes * 2);
  ncap = (ncap + HS_TILE - 1) / HS_TILE * HS_TILE;
  if (ncap < HS_TILE) ncap = HS_TILE;
  int4 *nd = nullptr, *ns = nullptr; uint32_t* nf = nullptr; int16_t* np = nullptr;
  CUDA_OK(h, cudaMalloc(&nd, (size_t)ncap * HS_D * sizeof(int4)));
  CUDA_OK(h, cudaMalloc(&ns, (size_t)ncap * HS_D * sizeof(int4)));
  CUDA_OK(h, cudaMalloc(&nf, (size_t)ncap * sizeof(uint32_t)));
  CUDA_OK(h, cudaMalloc(&np, (size_t)ncap * 64 * sizeof(int16_t)));
  int4* nt = nullptr;
  CUDA_OK(h, cudaMalloc(&nt, (size_t)ncap * sizeof(int4)));
  if (h->d_tot) cudaFree(h->d_tot);
  h->d_tot = nt;
  h->h_tot.resize((size_t)ncap, make_int4(0, 0, 0, 0));
  if (h->d_dyn) { cudaFree(h->d_dyn); cudaFree(h->d_stat); cudaFree(h->d_flags); cudaFree(h->d_pair); }
  h->d_dyn = nd; h->d_stat = ns; h->d_flags = nf; h->d_pair = np;
  h->h_dyn.resize((size_t)ncap * HS_D, make_int4(0, 0, 0, 0));
  h->h_stat.resize((size_t)ncap * HS_D, make_int4(0, 0, 0, 0));
  h->h_flags.resize((size_t)ncap, 0u);
  h->h_pair.resize((size_t)ncap * 64, (int16_t)0);
  h->cap_nodes = ncap;
  h->dirty_lo = 0;
  h->dirty_hi = ncap / HS_TILE - 1;
  return HS_OK;
}

inline size_t inv_idx(int node, int d) { return (size_t)node * HS_D + d; }   // node-major SoA (hs_kernels.cu)

void mark_dirty(hs_handle* h, int node) {
  int t = node / HS_TILE;
  h->dirty_lo = std::min(h->dirty_lo, t);
  h->dirty_hi = std::max(h->dirty_hi, t);
}

// every byte the batch path moves over PCIe is counted where it is copied (bench.py reports the measured figures)
inline cudaError_t up(hs_handle* h, void* dst, const void* src, size_t bytes) {
  h->bytes_h2d += bytes;
  return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->stream);
}
inline cudaError_t down(hs_handle* h, void* dst, const void* src, size_t bytes) {
  h->bytes_d2h += bytes;
  return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, h->stream);
}
inline cudaError_t down_sync(hs_handle* h, void* dst, const void* src, size_t bytes) {
  h->bytes_d2h += bytes;
  return cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost);
}

int flush_inventory(hs_handle* h) {
  if (h->mig_tab_uploaded != h->mig_geoms.size()) {      // the MIG geometry table: 16 x 272 B, rewritten when a new one appears
    if (!h->d_mig_tab) CUDA_OK(h, cudaMalloc(&h->d_mig_tab, 16 * sizeof(MigGeom)));
    std::vector<MigGeom> tab(16);
    memset(tab.data(), 0, 16 * sizeof(MigGeom));
    for (size_t g = 0; g < h->mig_geoms.size() && g < 16; g++) tab[g] = h->mig_geoms[g].dev;
    CUDA_OK(h, cudaMemcpy(h->d_mig_tab, tab.data(), 16 * sizeof(MigGeom), cudaMemcpyHostToDevice));
    h->mig_tab_uploaded = h->mig_geoms.size();
  }
  if (h->dirty_hi < h->dirty_lo) return HS_OK;
  size_t t0 = (size_t)h->dirty_lo, nt = (size_t)(h->dirty_hi - h->dirty_lo + 1);
  size_t o = t0 * HS_D * HS_TILE, n = nt * HS_D * HS_TILE;
  CUDA_OK(h, up(h, h->d_dyn + o, h->h_dyn.data() + o, n * sizeof(int4)));
  CUDA_OK(h, up(h, h->d_stat + o, h->h_stat.data() + o, n * sizeof(int4)));
  CUDA_OK(h, up(h, h->d_tot + t0 * HS_TILE, h->h_tot.data() + t0 * HS_TILE, nt * HS_TILE * sizeof(int4)));
  CUDA_OK(h, up(h, h->d_flags + t0 * HS_TILE, h->h_flags.data() + t0 * HS_TILE, nt * HS_TILE * sizeof(uint32_t)));
  CUDA_OK(h, up(h, h->d_pair + t0 * HS_TILE * 64, h->h_pair.data() + t0 * HS_TILE * 64, nt * HS_TILE * 64 * sizeof(int16_t)));
  CUDA_OK(h, cudaStreamSynchronize(h->stream));  // the pageable host vectors may be resized right after
  h->dirty_lo = 1 << 30;
  h->dirty_hi = -1;
  return HS_OK;
}

bool get_anno(const hs_pod* pod, const std::string& key, std::string& out) {
  if (key.empty()) return false;
  // Go map semantics: a later duplicate key replaces an earlier one
  bool found = false;
  for (uint32_t i = 0; i < pod->n_annotations; i++)
    if (pod->annotations[i].key && key == pod->annotations[i].key) {
      out = pod->annotations[i].value ? pod->annotations[i].value : "";
      found = true;
    }
  return found;
}

// masks over the interned device types for one (k.Type, selector annotations) combination
std::array<uint64_t, 4> entry_masks(hs_handle* h, const VendorInfo* v, const std::string& ktype, const hs_pod* pod) {
  if (h->mask_cache_types != h->types.size()) { h->mask_cache.clear(); h->mask_cache_types = h->types.size(); }
  std::string use, nouse, mode;
  bool hu = false, hn = false, hm = false;
  if (v) {
    hu = get_anno(pod, v->use_type, use);
    hn = get_anno(pod, v->nouse_type, nouse);
    if (v->word == "NVIDIA") hm = get_anno(pod, "nvidia.com/vgpu-mode", mode);
  }
  std::string key = ktype;
  key += '\x1f'; key += v ? v->word : "?";
  key += '\x1f'; if (hu) { key += '1'; key += use; }
  key += '\x1f'; if (hn) { key += '1'; key += nouse; }
  key += '\x1f'; if (hm) { key += '1'; key += mode; }
  auto it = h->mask_cache.find(key);
  if (it != h->mask_cache.end()) return it->second;
  std::array<uint64_t, 4> m = {0, 0, 0, 0};
  std::string kl = lower(ktype);
  for (size_t t = 0; t < h->types.size(); t++) {
    if (h->types_lower[t].find(kl) != std::string::npos) m[0] |= 1ull << t;   // getNodeResources, score.go:43-52
    if (h->types[t] == ktype) m[1] |= 1ull << t;                              // gpu_policy.go:203
    bool allow = true;
    if (v && !v->use_type.empty()) allow = checkType(hu ? &use : nullptr, hn ? &nouse : nullptr, h->types_upper[t]);
    if (v && v->neuron) allow = h->types[t].find(ktype) != std::string::npos;   // strings.Contains(dev.Type, k.Type), awsneuron/device.go:497
    if (allow) m[2] |= 1ull << t;
  }
  uint64_t modes = 0xffull;
  if (hm) {  // nvidia/device.go:495-498: ok && !strings.Contains(mode, d.Mode) → mismatch
    modes = 0;
    for (size_t i = 0; i < h->modes.size(); i++) if (mode.find(h->modes[i]) != std::string::npos) modes |= 1ull << i;
  }
  m[3] = modes;
  h->mask_cache[key] = m;
  return m;
}

// use/no-use UUID annotation → sorted (node<<8 | mask8) words appended to the pool
bool uuid_words(hs_handle* h, const std::string& list, std::vector<uint32_t>& pool, uint32_t& off, uint16_t& cnt) {
  std::map<uint32_t, uint32_t> per_node;
  for (auto& u0 : splitComma(list)) {
    std::string u = trimSpace(u0);
    auto it = h->uuid_index.find(u);
    if (it == h->uuid_index.end()) continue;
    for (uint32_t w : it->second) per_node[w >> 8] |= 1u << (w & 255u);
  }
  off = (uint32_t)pool.size();
  for (auto& kv : per_node) pool.push_back((kv.first << 8) | (kv.second & 255u));
  if (per_node.size() > 65535) return false;
  cnt = (uint16_t)per_node.size();
  return true;
}

struct Compiled {
  DevPod pod;
  std::vector<DevEntry> entries;
  OpMeta meta;
  uint32_t max_alloc = 0;
};

// Everything Filter derives from the pod before it looks at a node: Resourcereqs' output,
// the per-pod knobs (score.go:139-146,361-364; util.go:277-285; scheduler.go:641-652) and
// the selector annotations, flattened to integers.
// Row of (namespace, vendor) in this batch's device quota table, -1 when Fit has nothing to check there (no entry, or no
// limit set: FitQuota gates on LimitSet, quota.go:84,91).
int quota_row(hs_handle* h, const std::string& ns, int vid) {
  if (!h->any_quota_limit) return -1;
  const hs_handle::QuotaEnt* qe = h->quota_find(ns, vid);
  if (!qe || (!qe->set[0] && !qe->set[1])) return -1;
  for (size_t i = 0; i < h->quota_rows.size(); i++)
    if (h->quota_rows[i].second == vid && h->quota_rows[i].first == ns) return (int)i;
  h->quota_rows.push_back({ns, vid});
  return (int)h->quota_rows.size() - 1;
}
void quota_rows_for(hs_handle* h, const std::string& ns, DevPod& p) {
  p.q_idx[0] = quota_row(h, ns, HV_NVIDIA);
  p.q_idx[1] = quota_row(h, ns, HV_MLU);
}
// countPodDevices + addUsageLocked / rmUsageLocked (quota.go:97-184) on the host mirror, all vendors
void quota_track(hs_handle* h, const std::string& ns, int node, const int4* delta, int sign) {
  const NodeSlot& nsl = h->nodes[node];
  int64_t mem[32] = {0}, core[32] = {0};
  bool any[32] = {false};
  for (size_t d = 0; d < nsl.devs.size() && d < (size_t)HS_D; d++) {
    if (!(delta[d].x | delta[d].y | delta[d].z)) continue;
    const int v = nsl.devs[d].vendor_id & 31;
    mem[v] += delta[d].y; core[v] += delta[d].z; any[v] = true;
  }
  hs_handle::NsQuota* nq = nullptr;
  for (int v = 0; v < 32; v++) {
    if (!any[v]) continue;
    if (!nq) {
      if (h->last_quota && h->last_quota_ns == ns) nq = h->last_quota;          // streams are mostly one namespace
      else if (sign > 0) nq = &h->quota[ns];
      else { auto it = h->quota.find(ns); if (it == h->quota.end()) return; nq = &it->second; }
      h->last_quota = nq; h->last_quota_ns = ns;                                  // unordered_map nodes do not move
    }
    hs_handle::QuotaEnt& q = nq->v[v];
    if (sign > 0) {
      nq->present |= 1u << v;
      q.used[0] += mem[v]; q.used[1] += core[v];
    } else if ((nq->present >> v) & 1u) {
      q.used[0] = std::max<int64_t>(0, q.used[0] - mem[v]);
      q.used[1] = std::max<int64_t>(0, q.used[1] - core[v]);
    }
  }
}

int compile_pod(hs_handle* h, const hs_pod* pod, std::vector<uint32_t>& uuid_pool, Compiled& c) {
  memset(&c.pod, 0, sizeof c.pod);
  c.entries.clear();
  c.meta = OpMeta();
  c.max_alloc = 0;
  c.meta.uid = pod->uid ? pod->uid : "";
  c.meta.ns = pod->ns ? pod->ns : "";
  c.meta.n_ctr = (int)pod->n_containers;
  c.meta.n_init = (int)std::min(pod->n_init, pod->n_containers);
  bool any = false;
  for (uint32_t i = 0; i < pod->n_containers; i++) if (pod->containers[i].n_reqs > 0) any = true;
  if (!any) { c.meta.status = HS_PASSTHROUGH; return HS_PASSTHROUGH; }
  if (pod->n_containers > HS_MAX_C) { h->err = "pod has more containers than HS_MAX_CONTAINERS"; return HS_ERR_UNSUPPORTED; }

  std::string tmp;
  std::string gpuPolicy = h->gpu_policy, nodePolicy = h->node_policy;
  if (get_anno(pod, "hami.io/gpu-scheduler-policy", tmp)) gpuPolicy = tmp;
  if (get_anno(pod, "hami.io/node-scheduler-policy", tmp)) nodePolicy = tmp;
  bool numaBind = get_anno(pod, "nvidia.com/numa-bind", tmp) && parseBoolTrue(tmp);
  int64_t w[3] = {1, 1, 1};
  if (get_anno(pod, "hami.io/device-scoring-weights", tmp)) {
    std::string why;
    if (!parseWeights(tmp, w, why)) {      // the reference's text, util/device_scoring_weights.go:62-98
      h->err = "invalid hami.io/device-scoring-weights annotation " + quoteGo(tmp) + ": " + why;
      return HS_ERR_WEIGHTS;
    }
  }
  DevPod& p = c.pod;
  p.op = OP_FILTER;
  p.w_slot = (float)w[0]; p.w_core = (float)w[1]; p.w_mem = (float)w[2];
  if (nodePolicy == "spread") p.flags |= PF_NODE_SPREAD;
  if (nodePolicy == "binpack") p.flags |= PF_NODE_BINPACK;
  if (policyContains(gpuPolicy, "mutex")) p.flags |= PF_MUTEX;
  if (policyContains(gpuPolicy, "topology-aware")) p.flags |= PF_TOPOLOGY;
  if (numaBind) p.flags |= PF_NUMA_ASSERT;
  if (get_anno(pod, "huawei.com/vnpu-mode", tmp)) { c.meta.vnpu_mode = tmp; if (tmp == "hami-core") p.flags |= PF_ASCEND_HAMICORE; }
  // comparator, policy/gpu_policy.go:77-167
  if (gpuPolicy.find(',') != std::string::npos || gpuPolicy == "numa") {
    p.sort_mode = SORT_CHAIN;
    std::vector<int> chain;
    bool seen[4] = {false, false, false, false};
    for (auto& t0 : splitComma(gpuPolicy)) {
      std::string t = trimSpace(t0);
      int k = t == "binpack" ? CK_BINPACK : t == "spread" ? CK_SPREAD : t == "numa" ? CK_NUMA : CK_NONE;
      if (k != CK_NONE && !seen[k]) { chain.push_back(k); seen[k] = true; }
    }
    if (chain.empty()) chain.push_back(CK_SPREAD);
    if (numaBind && chain[0] != CK_NUMA) {
      std::vector<int> withNuma = {CK_NUMA};
      for (int k : chain) if (k != CK_NUMA) withNuma.push_back(k);
      chain = withNuma;
    }
    for (size_t i = 0; i < 3; i++) p.chain[i] = i < chain.size() ? (uint8_t)chain[i] : (uint8_t)CK_NONE;
  } else if (gpuPolicy == "mutex") {
    p.sort_mode = SORT_MUTEX;
  } else if (numaBind) {
    p.sort_mode = gpuPolicy == "binpack" ? SORT_NUMA_BINPACK : SORT_NUMA_SPREAD;
  } else {
    p.sort_mode = gpuPolicy == "binpack" ? SORT_BINPACK : SORT_SPREAD;
  }
  std::string vgpuMode;
  bool hasVgpuMode = get_anno(pod, "nvidia.com/vgpu-mode", vgpuMode);
  bool has_sgpu = false, sgpu_topo_default = false;

  p.n_ctr = (uint8_t)pod->n_containers;
  p.n_init = (uint8_t)c.meta.n_init;
  std::set<std::string> keyset, slotset;
  c.meta.ctr_begin.assign(pod->n_containers + 1, 0);
  for (uint32_t ci = 0; ci < pod->n_containers; ci++) {
    // ContainerDeviceRequests is a map: later duplicate keys replace earlier ones; canonical rule 4: sorted keys
    std::map<std::string, const hs_container_request*> m;
    for (uint32_t r = 0; r < pod->containers[ci].n_reqs; r++) {
      const hs_container_request* q = &pod->containers[ci].reqs[r];
      m[q->key ? q->key : ""] = q;
    }
    p.ctr_begin[ci] = (uint8_t)c.entries.size();
    c.meta.ctr_begin[ci] = (uint8_t)c.entries.size();
    c.meta.ctr_names.push_back(pod->containers[ci].name ? pod->containers[ci].name : "container-" + std::to_string(ci));
    for (auto& kv : m) {
      const hs_container_request* q = kv.second;
      if (c.entries.size() >= HS_MAX_E) { h->err = "pod has more request entries than HS_MAX_REQS_PER_POD"; return HS_ERR_UNSUPPORTED; }
      DevEntry e;
      memset(&e, 0, sizeof e);
      std::string ktype = q->type ? q->type : "";
      e.nums = q->nums; e.memreq = q->memreq; e.mempct = q->mempct; e.coresreq = q->coresreq;
      auto vit = h->vendor_by_word.find(ktype);
      const VendorInfo* v = vit == h->vendor_by_word.end() ? nullptr : &h->vendors[vit->second];
      e.vendor = v ? (uint16_t)v->gate : (uint16_t)HV_NONE;
      if (v && v->count_only) e.flags |= EF_COUNT_ONLY;
      if (v && v->vastai) e.flags |= EF_VASTAI;
      if (v && v->amd) e.flags |= EF_AMD;
      if (v && (v->mthreads || v->metax)) e.flags |= EF_MTHREADS;   // metax/device.go:156-166 is mthreads' customFilterRule
      if (v && v->metax) e.flags |= EF_METAX;
      if (v && v->neuron) e.flags |= EF_NEURON;
      if (v && v->whole_card) e.flags |= EF_WHOLE_CARD;
      if (v && v->kunlun) e.flags |= EF_KUNLUN;
      if (v && v->enflame) e.flags |= EF_ENFLAME;
      if (v && v->sgpu) {
        e.flags |= EF_SGPU;
        sgpu_topo_default = v->option != 0;
        has_sgpu = true;
      }
      if (v && v->vxpu) e.flags |= EF_VXPU;
      auto masks = entry_masks(h, v, ktype, pod);
      e.type_contains = masks[0]; e.type_equal = masks[1]; e.type_allow = masks[2]; e.mode_allow = (uint32_t)masks[3];
      if (v && v->ascend910) e.flags |= EF_ASCEND910;
      if (v && v->ascend && ktype == "Ascend910C") e.flags |= EF_ASCEND910C;
      if (!hasVgpuMode || vgpuMode.find("mig") != std::string::npos) e.flags |= EF_MIGMODE_OK;
      e.score_slot = -1;
      e.is_init = ci < (uint32_t)c.meta.n_init ? 1 : 0;
      if (v) {
        std::string list;
        if (get_anno(pod, v->use_uuid, list) && !trimSpace(list).empty()) {
          e.has_use = 1;
          if (!uuid_words(h, list, uuid_pool, e.use_off, e.use_cnt)) { h->err = "use-uuid list spans too many nodes"; return HS_ERR_UNSUPPORTED; }
        }
        std::string use2, nouse2;
        if (v->kunlun) {
          // kunlun's fitFn consults a second, fixed pair as well (kunlun/device.go:196-201): a card must pass both, so the
          // use lists intersect and the no-use lists unite
          if (get_anno(pod, "baidu.com/use-gpuuuid", use2) && !trimSpace(use2).empty()) {
            if (e.has_use) {
              std::set<std::string> a;
              for (auto& u : splitComma(list)) a.insert(trimSpace(u));
              std::string both;
              for (auto& u : splitComma(use2)) if (a.count(trimSpace(u))) { if (!both.empty()) both += ","; both += trimSpace(u); }
              if (both.empty()) both = "\x01";       // no card passes both lists (no UUID is this string)
              use2 = both;
            }
            e.has_use = 1;
            e.use_off = 0; e.use_cnt = 0;
            if (!uuid_words(h, use2, uuid_pool, e.use_off, e.use_cnt)) { h->err = "use-uuid list spans too many nodes"; return HS_ERR_UNSUPPORTED; }
          }
          if (get_anno(pod, "baidu.com/nouse-gpuuuid", nouse2) && !trimSpace(nouse2).empty()) {
            std::string first;
            if (get_anno(pod, v->nouse_uuid, first) && !trimSpace(first).empty()) nouse2 = first + "," + nouse2;
            e.nouse_off = 0; e.nouse_cnt = 0;
            if (!uuid_words(h, nouse2, uuid_pool, e.nouse_off, e.nouse_cnt)) { h->err = "no-use-uuid list spans too many nodes"; return HS_ERR_UNSUPPORTED; }
            nouse2 = "\x02";
          }
        }
        if (nouse2 != "\x02" && get_anno(pod, v->nouse_uuid, list) && !trimSpace(list).empty())
          if (!uuid_words(h, list, uuid_pool, e.nouse_off, e.nouse_cnt)) { h->err = "no-use-uuid list spans too many nodes"; return HS_ERR_UNSUPPORTED; }
      }
      keyset.insert(ktype);
      if (v && (v->ascend910 || v->metax || v->vastai || v->sgpu)) slotset.insert(ktype);   // vendors whose ScoreNode is not constant 0
      c.entries.push_back(e);
      c.meta.entries.push_back(EntryMeta{ktype, (int)ci, v ? v->id : -1});
      int nn = q->nums < 0 ? 0 : (q->nums > HS_D ? HS_D : q->nums);
      c.max_alloc += (uint32_t)nn;
    }
  }
  p.ctr_begin[pod->n_containers] = (uint8_t)c.entries.size();
  c.meta.ctr_begin[pod->n_containers] = (uint8_t)c.entries.size();
  p.n_entries = (uint8_t)c.entries.size();
  if (slotset.size() > HS_SCORE_SLOTS) { h->err = "pod names more Ascend910 types than HS_SCORE_SLOTS"; return HS_ERR_UNSUPPORTED; }
  if (c.max_alloc > HS_MAX_ALLOC) { h->err = "pod asks for more devices than HS_MAX_ALLOC"; return HS_ERR_UNSUPPORTED; }
  for (size_t i = 0; i < c.entries.size(); i++) {
    const std::string& t = c.meta.entries[i].type;
    c.entries[i].key_id = (uint8_t)std::distance(keyset.begin(), keyset.find(t));
    auto s = slotset.find(t);
    if (s != slotset.end()) c.entries[i].score_slot = (int8_t)std::distance(slotset.begin(), s);
  }
  if (has_sgpu) {       // metax sGPU reads three pod annotations (metax/protocol.go:33-37)
    std::string v;
    int cls = 0;
    if (get_anno(pod, "metax-tech.com/sgpu-app-class", v)) cls = v == "online" ? 1 : (v == "offline" ? 2 : 3);
    bool topo = sgpu_topo_default;
    if (get_anno(pod, "metax-tech.com/sgpu-topology-aware", v)) { bool b; if (parseBool(v, b)) topo = b; }
    p.sg_class = (uint8_t)(cls | (topo ? 4 : 0));
    int q = 0;
    if (get_anno(pod, "metax-tech.com/sgpu-qos-policy", v)) {
      q = 7;                                     // a value no device can carry …
      for (size_t i = 0; i < h->modes.size() && i < 7; i++) if (h->modes[i] == v) q = (int)i;   // … unless the table holds it
    }
    p.sg_qos = (uint8_t)q;
    c.meta.sg_class = cls; c.meta.sg_qos = q;
  }
  if (!c.entries.empty()) {
    p.e0 = c.entries[0];
    const DevEntry& e = c.entries[0];
    // Ascend910* (NetworkID groups + ScoreNode) rides the lane-per-device path too (fast_ascend_select); 910C's module
    // pairing stays with the general evaluator
    p.fast = (c.entries.size() == 1 && c.meta.n_init == 0 && e.nums >= 1 && e.nums <= HS_D && e.vendor != HV_NONE &&
              !(e.flags & EF_ASCEND910C)) ? 1 : 0;
    if (h->no_fast || (e.flags & (EF_COUNT_ONLY | EF_AMD | EF_MTHREADS | EF_WHOLE_CARD | EF_METAX | EF_NEURON | EF_KUNLUN | EF_ENFLAME | EF_SGPU))) p.fast = 0;
  }
  if (h->has_mig)       // dynamic MIG lives in the general evaluator: every NVIDIA request of a handle that holds a GPU in MIG mode
    for (const DevEntry& e : c.entries) {
      if (e.vendor != HV_NVIDIA) continue;
      p.fast = 0;
      if (!(e.flags & EF_MIGMODE_OK)) continue;      // the pod pins another mode: it never reaches a MIG GPU's placements
      if (e.nums > HS_D) { h->err = "more than 8 MIG instances in one request"; return HS_ERR_UNSUPPORTED; }
      if (p.flags & PF_TOPOLOGY) {
        // topology-aware Fit on a MIG GPU keeps visiting it until its placements run out and then chooses among all
        // (GPU, placement) candidates by pair score (nvidia/device.go:831-870): more candidates than the evaluator's lists hold
        h->err = "topology-aware requests on a handle with GPUs in MIG mode are outside the built scope";
        return HS_ERR_UNSUPPORTED;
      }
    }
  quota_rows_for(h, c.meta.ns, p);
  if (p.q_idx[0] >= 0 || p.q_idx[1] >= 0) p.fast = 0;   // the quota check lives in the general evaluator (hs_eval.cuh)
  c.meta.kind = OP_FILTER;
  c.meta.status = HS_OK;
  return HS_OK;
}

// NodeScore.ComputeDefaultScore (policy/node_policy.go:106-130) on the host mirror, float32 like the device
int host_node_score_bits(hs_handle* h, int node) {
  int32_t used = 0, usedCore = 0, usedMem = 0, total = 0, totalCore = 0, totalMem = 0;
  for (int d = 0; d < HS_D; d++) {
    const int4& dy = h->h_dyn[inv_idx(node, d)];
    const int4& st = h->h_stat[inv_idx(node, d)];
    used += dy.x; usedMem += dy.y; usedCore += dy.z;
    total += st.x; totalMem += st.y; totalCore += st.z;
  }
  float sc = 0.0f;
  if (!(total == 0 || totalCore == 0 || totalMem == 0)) {
    volatile float a = (float)used / (float)total;
    volatile float b = (float)usedCore / (float)totalCore;
    volatile float c = (float)usedMem / (float)totalMem;
    volatile float s1 = a + b;
    volatile float s2 = s1 + c;
    sc = 10.0f * s2;
  }
  int bits;
  memcpy(&bits, &sc, 4);
  return bits;
}

void apply_host_delta(hs_handle* h, int node, const int4* delta, int sign) {
  for (int d = 0; d < HS_D; d++) {
    int4& v = h->h_dyn[inv_idx(node, d)];
    v.x += sign * delta[d].x; v.y += sign * delta[d].y; v.z += sign * delta[d].z;
  }
  h->h_tot[node].y = host_node_score_bits(h, node);
}

// ---- dynamic MIG, host side ------------------------------------------------------------------------------------
// The device table entry for one MigProfiles list: selectMigCandidate's stable (MemoryMB, SliceCount) order, each
// candidate with the placements of the FIRST profile of its name (findMigProfile, mig_topology.go:42-49) in
// selectMigPlacement's order (descending start for "1g." / "3g." names, ascending otherwise; stable).
int intern_geom(hs_handle* h, const hs_mig_profile* profiles, uint32_t n, std::string& err) {
  if (n > 16) { err = "a GPU in MIG mode reports more than 16 MIG profiles"; return -1; }
  MigGeomHost g;
  for (uint32_t i = 0; i < n; i++) {
    MigGeomHost::Prof p;
    p.name = profiles[i].name ? profiles[i].name : "";
    p.memory_mb = profiles[i].memory_mb; p.core = profiles[i].core; p.slice_count = profiles[i].slice_count;
    if (profiles[i].n_placements > 8) { err = "a MIG profile reports more than 8 placements"; return -1; }
    if (p.core > 32767 || p.core < -32768) { err = "MIG profile core share outside int16"; return -1; }
    for (uint32_t j = 0; j < profiles[i].n_placements; j++) {
      const hs_mig_placement& q = profiles[i].placements[j];
      if (q.size == 0) { err = "a MIG placement of size 0 never conflicts: the reference's Fit would visit the GPU forever (nvidia/device.go:846-848)"; return -1; }
      if (q.start > 7 || q.size > 8 || q.start + q.size > 8) { err = "MIG placement outside the engine's 8-slice mask"; return -1; }
      p.placements.emplace_back(q.start, q.size);
    }
    g.profiles.push_back(p);
  }
  for (size_t k = 0; k < h->mig_geoms.size(); k++) {
    const auto& o = h->mig_geoms[k].profiles;
    if (o.size() != g.profiles.size()) continue;
    bool same = true;
    for (size_t i = 0; i < o.size() && same; i++)
      same = o[i].name == g.profiles[i].name && o[i].memory_mb == g.profiles[i].memory_mb && o[i].core == g.profiles[i].core &&
             o[i].slice_count == g.profiles[i].slice_count && o[i].placements == g.profiles[i].placements;
    if (same) return (int)k;
  }
  if (h->mig_geoms.size() >= 16) { err = "more than 15 distinct MIG geometries on one handle"; return -1; }
  memset(&g.dev, 0, sizeof g.dev);
  std::vector<int> order(g.profiles.size());
  for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    if (g.profiles[a].memory_mb != g.profiles[b].memory_mb) return g.profiles[a].memory_mb < g.profiles[b].memory_mb;
    return g.profiles[a].slice_count < g.profiles[b].slice_count;
  });
  g.dev.n_cand = (uint32_t)order.size();
  for (size_t c = 0; c < order.size(); c++) {
    const auto& cp = g.profiles[order[c]];
    int first = order[c];
    for (size_t i = 0; i < g.profiles.size(); i++) if (g.profiles[i].name == cp.name) { first = (int)i; break; }
    auto pl = g.profiles[first].placements;
    const bool high = cp.name.compare(0, 3, "1g.") == 0 || cp.name.compare(0, 3, "3g.") == 0;
    std::stable_sort(pl.begin(), pl.end(), [&](const std::pair<uint32_t, uint32_t>& a, const std::pair<uint32_t, uint32_t>& b) {
      return high ? a.first > b.first : a.first < b.first;
    });
    MigCand& mc = g.dev.cand[c];
    mc.mem = cp.memory_mb; mc.core = (int16_t)cp.core; mc.prof = (uint8_t)order[c];
    mc.n_place = (uint8_t)pl.size();
    for (size_t j = 0; j < pl.size(); j++) mc.place[j] = (uint8_t)((pl[j].first << 4) | pl[j].second);
  }
  h->mig_geoms.push_back(g);
  return (int)h->mig_geoms.size() - 1;
}

// strconv.Atoi on a profile name part (decimal, optional sign); false = not a number
static bool enf_atoi(const std::string& t, long long& out) {
  if (t.empty()) return false;
  size_t i = 0;
  bool neg = false;
  if (t[0] == '+' || t[0] == '-') { neg = t[0] == '-'; i = 1; }
  if (i >= t.size()) return false;
  unsigned long long v = 0;
  for (; i < t.size(); i++) {
    if (t[i] < '0' || t[i] > '9') return false;
    if (v > (0x7fffffffffffffffULL - (unsigned)(t[i] - '0')) / 10) return false;      // out of int64 range: Atoi errors
    v = v * 10 + (unsigned)(t[i] - '0');
  }
  out = neg ? -(long long)v : (long long)v;
  return true;
}
// parseProfile, enflame/device.go:621-643: "<slices>g.<GB>gb"
static void enf_parse_profile(const std::string& name, long long& size, long long& mem_gb) {
  size = 0; mem_gb = 0;
  std::string n = trimSpace(name);
  for (auto& c : n) if (c >= 'A' && c <= 'Z') c = (char)(c + 32);
  std::vector<std::string> parts;
  size_t pos = 0;
  for (;;) { size_t q = n.find('.', pos); if (q == std::string::npos) { parts.push_back(n.substr(pos)); break; } parts.push_back(n.substr(pos, q - pos)); pos = q + 1; }
  if (parts.size() < 2) return;
  std::string sp = trimSpace(parts[0]);
  if (!sp.empty() && sp.back() == 'g') sp.pop_back();
  long long sz;
  if (!enf_atoi(sp, sz)) return;
  size = sz;
  std::string mp = trimSpace(parts[1]);
  if (mp.size() >= 2 && mp.compare(mp.size() - 2, 2, "gb") == 0) mp.erase(mp.size() - 2);
  long long mg;
  if (!enf_atoi(mp, mg)) return;
  mem_gb = mg;
}
// The node's candidate table (collectDRSProfiles): every profile name the node's vGCU devices advertise that parses to a
// positive slice count and memory, a later device's id winning, core share = ceil(slices * 100 / maxSlice), sorted by
// (slices, GB) — equal ones in ascending name (the reference iterates a Go map there).
int intern_enflame(hs_handle* h, const std::vector<DevSlot*>& devs, std::string& err) {
  long long maxSlice = 0;
  for (auto* d : devs) if (d->enf_max_slice > maxSlice) maxSlice = d->enf_max_slice;
  if (maxSlice <= 0)
    for (auto* d : devs)
      for (auto& kv : d->enf_profiles) { long long s, m; enf_parse_profile(kv.first, s, m); if (s > maxSlice) maxSlice = s; }
  MigGeomHost g;
  g.enflame = true;
  memset(&g.dev, 0, sizeof g.dev);
  if (maxSlice > 0) {
    std::map<std::string, MigGeomHost::Enf> seen;
    for (auto* d : devs)
      for (auto& kv : d->enf_profiles) {
        long long s, m;
        enf_parse_profile(kv.first, s, m);
        if (s <= 0 || m <= 0) continue;
        seen[kv.first] = MigGeomHost::Enf{kv.first, kv.second, s, m, (long long)std::ceil((double)s * 100 / (double)maxSlice)};
      }
    for (auto& kv : seen) g.enf.push_back(kv.second);
    std::stable_sort(g.enf.begin(), g.enf.end(), [](const MigGeomHost::Enf& a, const MigGeomHost::Enf& b) {
      if (a.size == b.size) return a.mem_gb < b.mem_gb;
      return a.size < b.size;
    });
  }
  if (g.enf.size() > 16) { err = "an enflame node advertises more than 16 usable DRS profiles"; return -1; }
  for (size_t k = 0; k < h->mig_geoms.size(); k++) {
    const auto& o = h->mig_geoms[k];
    if (!o.enflame || o.enf.size() != g.enf.size()) continue;
    bool same = true;
    for (size_t i = 0; i < o.enf.size() && same; i++)
      same = o.enf[i].name == g.enf[i].name && o.enf[i].id == g.enf[i].id && o.enf[i].core == g.enf[i].core;
    if (same) return (int)k;
  }
  if (h->mig_geoms.size() >= 16) { err = "more than 15 distinct MIG geometries / enflame profile tables on one handle"; return -1; }
  g.dev.n_cand = (uint32_t)g.enf.size();
  for (size_t c = 0; c < g.enf.size(); c++) {
    const auto& p = g.enf[c];
    MigCand& mc = g.dev.cand[c];
    // the reference turns these into "1/n ModeNotFit" when selected (enflame/device.go:393-408)
    if (p.core > 32767) { err = "enflame profile core share outside the engine's int16 (slice count far above maxSlice)"; return -1; }
    const bool valid = p.size <= 2147483647LL && p.mem_gb <= 2147483647LL / 1024;
    mc.mem = valid ? (int32_t)(p.mem_gb * 1024) : -1;
    mc.core = (int16_t)(p.core > 32767 ? 32767 : p.core);
    mc.prof = (uint8_t)c;
    const int32_t sz = (int32_t)std::min<long long>(p.size, 2147483647LL), gb = (int32_t)std::min<long long>(p.mem_gb, 2147483647LL);
    for (int b = 0; b < 4; b++) { mc.place[b] = (uint8_t)((uint32_t)sz >> (8 * b)); mc.place[4 + b] = (uint8_t)((uint32_t)gb >> (8 * b)); }
  }
  h->mig_geoms.push_back(g);
  return (int)h->mig_geoms.size() - 1;
}

inline uint32_t mig_res_mask(uint32_t start, uint32_t size) {
  uint32_t m = 0;
  for (uint32_t b = start; b < 8 && b - start < size; b++) m |= 1u << b;
  return m;
}

// What ONE cached pod does to the MIG state of its node (the pod's share of getNodesUsage, scheduler.go:757-842):
// occ[d] = slices its reservations cover on device d, bad[d] = 1 when it makes the device fail closed.
void mig_pod_effect(const hs_handle* h, const NodeSlot& ns, const PodRec& pr, uint8_t* occ, uint8_t* bad) {
  for (int d = 0; d < HS_D; d++) { occ[d] = 0; bad[d] = 0; }
  const size_t nd = std::min(ns.devs.size(), (size_t)HS_D);
  std::vector<char> taken(pr.mig.size(), 0);
  for (size_t d = 0; d < nd; d++) {
    const DevSlot& ds = ns.devs[d];
    for (int k = 0; k < (int)pr.n_ent[d]; k++) {
      size_t r = 0;
      while (r < pr.mig.size() && (taken[r] || pr.mig[r].uuid != ds.uuid)) r++;
      if (r < pr.mig.size()) {
        if (ds.mode == "hami-core") { bad[d] = 1; continue; }         // :788-792 (the reservation is not consumed)
        taken[r] = 1;
        occ[d] |= (uint8_t)mig_res_mask(pr.mig[r].start, pr.mig[r].size);
        continue;
      }
      if (ds.mode == "mig") bad[d] = 1;                                // :799-802
    }
  }
  for (size_t r = 0; r < pr.mig.size(); r++) {                         // :810-841, reservations nothing consumed
    if (taken[r]) continue;
    bool matched = false;
    for (size_t d = 0; d < nd; d++) {
      if (ns.devs[d].uuid != pr.mig[r].uuid) continue;
      matched = true;
      if (ns.devs[d].mode != "mig") bad[d] = 1;
      else occ[d] |= (uint8_t)mig_res_mask(pr.mig[r].start, pr.mig[r].size);
      break;
    }
    if (!matched)
      for (size_t d = 0; d < nd; d++) if (ns.devs[d].mode == "mig") bad[d] = 1;
  }
  (void)h;
}

// The node's MIG state from the pod cache, leaving out the pods in `without`: occ / bad per device slot.  Returns
// false when the node has no GPU in MIG mode and no cached pod with reservations touches it (nothing to keep).
bool mig_node_state(const hs_handle* h, int slot, const std::unordered_set<std::string>* without, uint8_t* occ, int* bad) {
  for (int d = 0; d < HS_D; d++) { occ[d] = 0; bad[d] = 0; }
  const NodeSlot& ns = h->nodes[slot];
  if (!ns.has_mig && h->n_mig_pods == 0) return false;
  bool any = ns.has_mig;
  for (const auto& kv : h->pods) {
    if (kv.second.node != slot) continue;
    if (!ns.has_mig && kv.second.mig.empty()) continue;
    if (without && without->count(kv.first)) continue;
    uint8_t o[HS_D], b[HS_D];
    mig_pod_effect(h, ns, kv.second, o, b);
    for (int d = 0; d < HS_D; d++) { occ[d] |= o[d]; bad[d] += b[d]; }
    any = true;
  }
  return any;
}

inline uint32_t mig_dflags(uint32_t df, uint32_t occ, int bad) {
  bad = bad > 255 ? 255 : bad;
  df = (df & 0x0000ffffu) | (occ << 16) | ((uint32_t)bad << 24);
  return (bad == 0 && (df & DF_HEALTH_REG)) ? (df | DF_HEALTH) : (df & ~DF_HEALTH);
}

// A pod WITHOUT reservations joins (+1) or leaves (-1) the cache: it only moves the fail-closed count of the MIG GPUs it
// has entries on (mig_pod_effect with an empty reservation list), so the mirror follows without walking the pod cache —
// the case of every commit of a batch.
void mig_apply_orphan(hs_handle* h, int slot, const PodRec& pr, int sign) {
  const NodeSlot& ns = h->nodes[slot];
  for (size_t d = 0; d < ns.devs.size() && d < (size_t)HS_D; d++) {
    if (pr.n_ent[d] == 0 || ns.devs[d].mode != "mig") continue;
    int4& v = h->h_dyn[inv_idx(slot, (int)d)];
    const uint32_t df = (uint32_t)v.w;
    int bad = (int)DF_BAD(df) + sign;
    bad = bad < 0 ? 0 : bad;
    v.w = (int)mig_dflags(df, DF_OCC(df), bad);
  }
}

// metax sGPU: a pod of class `cls` (1 online, 2 offline) joins (+1) or leaves (-1) the PodInfos of the sGPU devices it has
// entries on — once per cached entry (scheduler.go:786)
void sgpu_apply_class(hs_handle* h, int slot, const uint8_t* n_ent, int cls, int sign) {
  if (cls != 1 && cls != 2) return;
  const int sh = cls == 1 ? 16 : 22;
  for (int d = 0; d < HS_D; d++) {
    int4& v = h->h_dyn[inv_idx(slot, d)];
    if (!((uint32_t)v.w & DF_SGPU) || n_ent[d] == 0) continue;
    int c = (int)(((uint32_t)v.w >> sh) & 63u) + sign * (int)n_ent[d];
    c = c < 0 ? 0 : (c > 63 ? 63 : c);
    v.w = (int)(((uint32_t)v.w & ~(63u << sh)) | ((uint32_t)c << sh));
  }
}

// Bring the node's dflags (host mirror) in line with the pod cache; true when something changed.
bool mig_refresh(hs_handle* h, int slot) {
  NodeSlot& ns = h->nodes[slot];
  uint8_t occ[HS_D]; int bad[HS_D];
  const bool any = mig_node_state(h, slot, nullptr, occ, bad);
  if (!any && !ns.mig_touched) return false;
  ns.mig_touched = any;
  bool changed = false;
  for (int d = 0; d < HS_D; d++) {
    int4& v = h->h_dyn[inv_idx(slot, d)];
    if ((uint32_t)v.w & DF_SGPU) continue;       // a metax sGPU device keeps its own bookkeeping in those bits
    const uint32_t nf = mig_dflags((uint32_t)v.w, occ[d], bad[d]);
    if (nf != (uint32_t)v.w) { v.w = (int)nf; changed = true; }
  }
  return changed;
}

// encoding/json string: quotes, backslash, control characters and the HTML trio escaped
std::string json_string(const char* v) {
  std::string o = "\"";
  for (const unsigned char* p = (const unsigned char*)(v ? v : ""); *p; p++) {
    if (*p == '"' || *p == '\\') { o += '\\'; o += (char)*p; }
    else if (*p < 0x20 || *p == '<' || *p == '>' || *p == '&') { char b[8]; snprintf(b, sizeof b, "\\u%04x", *p); o += b; }
    else o += (char)*p;
  }
  return o + "\"";
}

std::string encode_pod_single(const std::vector<std::vector<const hs_alloc_entry*>>& pd, const NodeSlot* type_from = nullptr) {
  // EncodePodSingleDevice / EncodeContainerDevices, devices.go:451-484.  type_from: the vendor's picks carry the
  // DEVICE's type string instead of k.Type (amd/device.go:341-346)
  std::string res;
  char buf[64];
  for (auto& ctr : pd) {
    for (auto* e : ctr) {
      const char* type = e->type_key;
      if (type_from)
        for (const auto& d : type_from->devs) if (d.uuid == e->uuid) { type = d.type.c_str(); break; }
      res += e->uuid; res += ','; res += type;
      snprintf(buf, sizeof buf, ",%d,%d:", e->usedmem, e->usedcores);
      res += buf;
    }
    res += ';';
  }
  return res;
}

const std::string& fail_string_slow(hs_handle* h, const DevFail& f);
inline const std::string& fail_string(hs_handle* h, const DevFail& f) {
  // a no-fit result names every node; at ~70 % occupancy neighbouring nodes differ, so the lookup runs once per
  // node: a direct-mapped table of reason records in front of the map (references into the map stay valid)
  const uint32_t hsh = (f.kind_cards * 0x9E3779B1u) ^ (f.c0 * 0x85EBCA6Bu) ^ (f.c1 * 0xC2B2AE35u);
  hs_handle::FailTab& t = h->fail_tab[(hsh >> 16) & (hs_handle::kFailTab - 1)];
  if (t.s && t.k0 == f.kind_cards && t.k1 == f.c0 && t.k2 == f.c1 && t.k3 == f.pad) return *t.s;
  const std::string& s = fail_string_slow(h, f);
  t.k0 = f.kind_cards; t.k1 = f.c0; t.k2 = f.c1; t.k3 = f.pad; t.s = &s;
  return s;
}
const std::string& fail_string_slow(hs_handle* h, const DevFail& f) {
  std::array<uint32_t, 4> key = {f.kind_cards, f.c0, f.c1, f.pad};
  auto it = h->reason_cache.find(key);
  if (it != h->reason_cache.end()) return it->second;
  std::string s;
  uint32_t kind = f.kind_cards & 15u, cards = f.kind_cards >> 4;
  if (kind == FK_MIG_ADD_FAILED) {        // score.go:125-128 + nvidia/device.go:649
    std::string uuid;
    if (f.c1 < h->nodes.size() && cards < h->nodes[f.c1].devs.size()) uuid = h->nodes[f.c1].devs[cards].uuid;
    s = "AddResourceUsage failed for device " + uuid + ": MIG profile and placement allocation failed";
  }
  else if (kind == FK_NODE_INSUFFICIENT) s = "NodeInsufficientDevice";
  else if (kind == FK_TYPE_NOT_FOUND) s = "Device type not found";
  else if (kind == FK_HIST || kind == FK_HIST_ZALLOC) {
    // common.GenReason, common/common.go:46-53
    std::vector<std::string> parts;
    char buf[96];
    for (int r = 0; r < HS_R_COUNT; r++) {
      uint32_t cnt = r < 8 ? (f.c0 >> (4 * r)) & 15u : (r < 16 ? (f.c1 >> (4 * (r - 8))) & 15u : (f.pad >> (4 * (r - 16))) & 15u);
      if (!cnt && !(kind == FK_HIST_ZALLOC && r == 12)) continue;      // a key the reference wrote with 0 is still printed
      snprintf(buf, sizeof buf, "%u/%u %s", cnt, cards, kReasonNames[r]);
      parts.push_back(buf);
    }
    std::sort(parts.begin(), parts.end());
    for (size_t i = 0; i < parts.size(); i++) { if (i) s += ", "; s += parts[i]; }
  }
  return h->reason_cache[key] = s;
}

int build_kernel_args(hs_handle* h, KernelArgs& ka, int n_ops, uint32_t alloc_cap) {
  memset(&ka, 0, sizeof ka);
  ka.dyn = h->d_dyn; ka.stat = h->d_stat; ka.node_flags = h->d_flags; ka.node_tot = h->d_tot; ka.pair = h->d_pair;
  ka.mig_tab = h->mig_geoms.empty() ? nullptr : h->d_mig_tab;
  ka.no_resident = h->no_resident ? 1 : 0;
  ka.n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  ka.cand_pos = h->have_cand ? h->b_candpos.d : nullptr;
  ka.pods = h->b_pods.d; ka.entries = h->b_entries.d; ka.uuid_pool = h->b_uuid.d; ka.rel_pool = h->b_rel.d;
  ka.n_ops = n_ops;
  ka.decisions = h->b_dec.d; ka.allocs = h->b_alloc.d; ka.alloc_cap = alloc_cap;
  ka.commit_delta = h->b_delta.d; ka.sync = h->b_best.d;
  ka.fail = h->b_fail.d; ka.fail_cap = h->fail_cap_now;
  ka.counters = h->b_counters.d;
  ka.pos_base = (int32_t)h->pos_base;
  ka.timing = h->timing ? h->b_timing.d : nullptr;
  ka.trace = h->trace ? h->b_trace.d : nullptr;
  ka.trace_ops = hs_handle::kTraceOps; ka.trace_stride = hs_handle::kTraceStride;
  ka.quota = h->quota_rows.empty() ? nullptr : h->b_quota.d;
  ka.sync2 = h->b_best.d + h->sync2_base;
  ka.world = h->xgpu_connected ? h->shard_world : 1;
  ka.rank = h->shard_rank;
  for (int p = 0; p < 8; p++) ka.mbox_peer[p] = h->mbox_peer[p];
  ka.xresult = h->d_mbox ? h->d_mbox + 32 : nullptr;
  ka.xerror = h->d_mbox ? reinterpret_cast<uint32_t*>(h->d_mbox + 36) : nullptr;
  ka.epoch_base = h->epoch_base;
  // windows: one GPU, the whole inventory as candidates, 256-thread CTAs (a quarter warp per tracked node)
  const bool win = h->windows && h->w_skey && h->shard_world <= 1 && !h->force_shard && !h->have_cand && h->block == 256 &&
                   ka.n_nodes <= h->w_cap_nodes && ka.n_nodes / HS_TILE <= HS_WMAX_TILES;
  ka.w_skey = win ? h->w_skey : nullptr;
  ka.w_rmask = h->w_rmask; ka.w_head = h->w_head; ka.w_rank = h->w_rank; ka.w_stats = h->w_stats;
  return HS_OK;
}

// Run the device ops [0, n) staged in the b_* host mirrors; drains the fail buffer and resumes
// until every op ran.  On return b_dec.h / b_alloc.h / b_delta.h hold the results.
// The device side of a batch in three parts, so that the host can compile the next chunk of ops while the device runs
// the current one (run_ops): device_begin sizes and clears the buffers for up to max_ops ops, device_range uploads the
// records of ops [first, last) and runs them (draining the reason buffer and resuming as often as needed),
// device_end brings the allocation rows and the collapsed deltas back.
constexpr int kSmallDirectOps = 4;
int device_begin(hs_handle* h, int max_ops, uint32_t alloc_bound) {
  int rc = flush_inventory(h);
  if (rc != HS_OK) return rc;
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  // A no-fit decision leaves 16 B per node; a launch stops when the buffer is full and the host drains it.  Up to 32 MB
  // of it, so a batch normally runs in ONE launch; the fused multi-GPU path keeps the fixed size, because every rank
  // must stop at the same op and the shards differ in size.
  h->fail_cap_now = h->fail_cap;
  if (!h->fail_cap_fixed && !h->xgpu_connected)
    h->fail_cap_now = (uint32_t)std::min<size_t>((size_t)std::max(max_ops, 1),
                                                 std::max<size_t>(h->fail_cap, (32u << 20) / ((size_t)std::max(n_nodes, HS_TILE) * sizeof(DevFail))));
  if (!h->b_dec.ensure(max_ops) || !h->b_alloc.ensure(std::max<uint32_t>(alloc_bound, 1)) || !h->b_delta.ensure((size_t)max_ops * HS_D) ||
      !h->b_best.ensure((size_t)max_ops * 2) || !h->b_counters.ensure(8) || !h->b_fail.ensure((size_t)h->fail_cap_now * std::max(n_nodes, HS_TILE))) {
    h->err = "out of device memory for batch buffers";
    return HS_ERR_CUDA;
  }
  if (h->windows && h->w_cap_nodes < n_nodes && n_nodes / HS_TILE <= HS_WMAX_TILES) {
    if (h->w_skey) { cudaFree(h->w_skey); cudaFree(h->w_rmask); cudaFree(h->w_head); cudaFree(h->w_rank); h->w_skey = nullptr; }
    const size_t cap = (size_t)std::max(n_nodes, HS_TILE), tiles = cap / HS_TILE;
    if (!h->w_stats) { CUDA_OK(h, cudaMalloc(&h->w_stats, 8 * sizeof(uint32_t))); CUDA_OK(h, cudaMemset(h->w_stats, 0, 8 * sizeof(uint32_t))); }
    CUDA_OK(h, cudaMalloc(&h->w_skey, cap * sizeof(unsigned long long)));
    CUDA_OK(h, cudaMalloc(&h->w_rmask, tiles * HS_WIN * sizeof(uint32_t)));
    CUDA_OK(h, cudaMalloc(&h->w_head, tiles * HS_WIN * sizeof(unsigned long long)));
    CUDA_OK(h, cudaMalloc(&h->w_rank, cap));
    h->w_cap_nodes = (int)cap;
  }
  h->sync2_base = max_ops;
  if (h->have_cand) CUDA_OK(h, up(h, h->b_candpos.d, h->b_candpos.h, (size_t)n_nodes * sizeof(int32_t)));
  // A handful of ops (one hs_filter): the kernel writes decisions, deltas, rows and counters straight into the pinned
  // host mirrors (page-locked memory is device-addressable): no result copies, no device-side clearing of them
  h->small_direct = max_ops <= kSmallDirectOps && h->shard_world <= 1 && !h->force_shard && !h->no_small_direct;
  if (h->small_direct) memset(h->b_dec.h, 0xff, (size_t)max_ops * sizeof(DevDecision));
  else CUDA_OK(h, cudaMemsetAsync(h->b_dec.d, 0xff, (size_t)max_ops * sizeof(DevDecision), h->stream));
  CUDA_OK(h, cudaMemsetAsync(h->b_best.d, 0, (size_t)max_ops * 2 * sizeof(OpSync), h->stream));   // [0,n): arg-best slots, [n,2n): quota barriers
  if (!h->quota_rows.empty())
    CUDA_OK(h, up(h, h->b_quota.d, h->b_quota.h, h->quota_rows.size() * 4 * sizeof(long long)));
  const int block = h->block;
  if (h->maxb_block != block) {
    h->maxb_cache[0] = hs_decide_max_blocks(block, h->device, 0);
    h->maxb_cache[1] = (h->fast_kernel && block == 256) ? hs_decide_max_blocks(block, h->device, 1) : 0;
    h->maxb_block = block;
  }
  const int maxb = h->maxb_cache[0];
  if (maxb <= 0) { h->err = "hs_decide_kernel cannot be launched cooperatively on this device"; return HS_ERR_CUDA; }
  h->grid_now = std::min(std::min(maxb, 1024), std::max(1, (n_nodes * HS_D + block - 1) / block));   // 8 lanes per node when it fits
  h->grid_fast = h->maxb_cache[1] > 0 ? std::min(std::min(h->maxb_cache[1], 1024), std::max(1, (n_nodes * HS_D + block - 1) / block)) : 0;
  h->fail_segs.clear();
  h->op_fail_row.assign(max_ops, nullptr);
  h->last_ms = 0;
  return HS_OK;
}

struct StageMark { size_t pods = 0, entries = 0, uuid = 0, rel = 0; };

// May ops [first, last) run on the instantiation without the general evaluator?  Only common-shape Filters and releases.
// Does any op of [first, last) read the entry pool?  (A pod's first entry rides in its own record.)
bool range_needs_entries(const hs_handle* h, int first, int last) {
  for (int i = first; i < last; i++)
    if (h->b_pods.h[i].op == OP_FILTER && h->b_pods.h[i].n_entries > 1) return true;
  return false;
}
bool range_is_fast(const hs_handle* h, int first, int last) {
  if (!h->grid_fast) return false;
  for (int i = first; i < last; i++)
    if (h->b_pods.h[i].op == OP_FILTER && !h->b_pods.h[i].fast) return false;
  return true;
}   // how much of each staging array is on the device

// with_rows: the collapsed deltas of these ops and their allocation rows [alloc_from, alloc_to) come back together with
// the decisions (otherwise the caller fetches rows and deltas for the whole batch afterwards, device_end).
int device_range(hs_handle* h, int first, int last, const StageMark& from, const StageMark& to, uint32_t alloc_cap,
                 bool with_rows = false, uint32_t alloc_from = 0, uint32_t alloc_to = 0) {
  cudaSetDevice(h->device);               // may run on a helper thread
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  const int n_ops = last;
  const int block = h->block, grid = h->grid_now;
  if (to.pods > from.pods) CUDA_OK(h, up(h, h->b_pods.d + from.pods, h->b_pods.h + from.pods, (to.pods - from.pods) * sizeof(DevPod)));
  if (to.entries > from.entries && range_needs_entries(h, first, last))
    CUDA_OK(h, up(h, h->b_entries.d + from.entries, h->b_entries.h + from.entries, (to.entries - from.entries) * sizeof(DevEntry)));
  if (to.uuid > from.uuid) CUDA_OK(h, up(h, h->b_uuid.d + from.uuid, h->b_uuid.h + from.uuid, (to.uuid - from.uuid) * sizeof(uint32_t)));
  if (to.rel > from.rel) CUDA_OK(h, up(h, h->b_rel.d + from.rel, h->b_rel.h + from.rel, (to.rel - from.rel) * sizeof(int4)));
  const bool direct = h->small_direct && with_rows;
  if (direct) memset(h->b_counters.h, 0, 8 * sizeof(uint32_t));
  else CUDA_OK(h, cudaMemsetAsync(h->b_counters.d, 0, 8 * sizeof(uint32_t), h->stream));
  int done = first;
  while (done < n_ops) {
    KernelArgs ka;
    build_kernel_args(h, ka, n_ops, alloc_cap);
    ka.op_begin = done;
    if (direct) { ka.decisions = h->b_dec.h; ka.commit_delta = h->b_delta.h; ka.allocs = h->b_alloc.h; ka.counters = h->b_counters.h; }
    CUDA_OK(h, cudaEventRecord(h->ev0, h->stream));
    const bool fk = range_is_fast(h, done, n_ops);
    cudaError_t le = (cudaError_t)hs_launch_decide(&ka, fk ? h->grid_fast : grid, block, h->stream, fk ? 1 : 0);
    if (le != cudaSuccess) { h->err = std::string("hs_decide_kernel launch: ") + cudaGetErrorString(le); return HS_ERR_CUDA; }
    h->launches++;
    CUDA_OK(h, cudaEventRecord(h->ev1, h->stream));
    // everything the launch produced comes back behind it on the stream, ONE synchronisation for all of it: counters,
    // the decisions of every op it may have reached (a launch normally runs to the end of its range), their collapsed
    // deltas and, for a pipelined chunk, its allocation rows
    if (!direct) {
    CUDA_OK(h, down(h, h->b_counters.h, h->b_counters.d, 8 * sizeof(uint32_t)));
    CUDA_OK(h, down(h, h->b_dec.h + done, h->b_dec.d + done, (size_t)(n_ops - done) * sizeof(DevDecision)));
    }
    if (with_rows && !direct) {
      CUDA_OK(h, down(h, h->b_delta.h + (size_t)done * HS_D, h->b_delta.d + (size_t)done * HS_D, (size_t)(n_ops - done) * HS_D * sizeof(int4)));
      if (alloc_to > alloc_from) CUDA_OK(h, down(h, h->b_alloc.h + alloc_from, h->b_alloc.d + alloc_from, (size_t)(alloc_to - alloc_from) * sizeof(DevAlloc)));
    }
    CUDA_OK(h, cudaStreamSynchronize(h->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    h->last_ms += ms;
    uint32_t ran = h->b_counters.h[2] - (uint32_t)done, nfail = h->b_counters.h[1];
    h->epoch_base += h->b_counters.h[5];
    if (h->b_counters.h[6]) { h->err = "fused multi-GPU arg-best: a peer rank never arrived (ranks must submit the same ops)"; return HS_ERR_CUDA; }
    const DevFail* seg = nullptr;
    if (nfail) {
      CUDA_OK(h, down(h, h->b_fail.h, h->b_fail.d, (size_t)nfail * n_nodes * sizeof(DevFail)));
      CUDA_OK(h, cudaStreamSynchronize(h->stream));
      h->fail_segs.emplace_back(h->b_fail.h, h->b_fail.h + (size_t)nfail * n_nodes);
      seg = h->fail_segs.back().data();
    }
    if (ran == 0 && nfail == 0 && n_ops - done > 0) { h->err = "decide kernel made no progress"; return HS_ERR_CUDA; }
    for (uint32_t i = 0; i < ran; i++) {
      DevDecision& d = h->b_dec.h[done + i];
      if (d.status == 1 && d.fail_slot < nfail) h->op_fail_row[done + i] = seg + (size_t)d.fail_slot * n_nodes;
    }
    done += (int)ran;
    if (done < n_ops) {
      // resume: keep the alloc cursor, clear the fail cursor
      h->b_counters.h[1] = 0; h->b_counters.h[2] = 0; h->b_counters.h[3] = 0; h->b_counters.h[5] = 0;
      CUDA_OK(h, cudaMemsetAsync(h->b_best.d + done, 0, sizeof(OpSync), h->stream));  // op `done` runs again
      CUDA_OK(h, cudaMemsetAsync(h->b_best.d + h->sync2_base + done, 0, sizeof(OpSync), h->stream));   // … and so may a window that starts there
      CUDA_OK(h, up(h, h->b_counters.d, h->b_counters.h, 8 * sizeof(uint32_t)));
    }
  }
  // (allocation rows sit at each op's own alloc_base, a host-made prefix sum, not behind a device cursor)
  return HS_OK;
}

// ---- NCCL, bound at run time (whichever libnccl.so.2 the process already has: torch's or the system's) ----
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, struct NcclId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
struct NcclId { char internal[128]; };
NcclApi g_nccl;
bool nccl_load(std::string& err) {
  if (g_nccl.lib) return true;
  void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) { err = std::string("dlopen libnccl.so.2: ") + dlerror(); return false; }
  g_nccl.GetUniqueId = (int (*)(void*))dlsym(lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(lib, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(lib, "ncclAllReduce");
  g_nccl.CommDestroy = (int (*)(void*))dlsym(lib, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce) { err = "libnccl.so.2 lacks the expected symbols"; return false; }
  g_nccl.lib = lib;
  return true;
}

// Sharded run of the staged device ops: per FILTER op sweep → allreduce(max) → decide → explain, all enqueued
// on the engine's stream with no host synchronisation in between; one sync at the end.
int run_device_ops_sharded(hs_handle* h, int n_ops, size_t n_entries, size_t n_uuid, size_t n_rel, uint32_t alloc_cap) {
  int rc = flush_inventory(h);
  if (rc != HS_OK) return rc;
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  const int nn = std::max(n_nodes, HS_TILE);
  // no resume in this mode: size the FailedNodes buffer for the batch (bounded at 64 MB)
  const uint32_t shard_fail_cap = (uint32_t)std::min<size_t>((size_t)n_ops, std::max<size_t>(h->fail_cap, (64u << 20) / ((size_t)nn * sizeof(DevFail))));
  if (!h->b_dec.ensure(n_ops) || !h->b_alloc.ensure(std::max<uint32_t>(alloc_cap, 1)) || !h->b_delta.ensure((size_t)n_ops * HS_D) ||
      !h->b_best.ensure(n_ops) || !h->b_counters.ensure(8) || !h->b_fail.ensure((size_t)shard_fail_cap * nn)) {
    h->err = "out of device memory for batch buffers";
    return HS_ERR_CUDA;
  }
  CUDA_OK(h, up(h, h->b_pods.d, h->b_pods.h, (size_t)n_ops * sizeof(DevPod)));
  if (n_entries) CUDA_OK(h, up(h, h->b_entries.d, h->b_entries.h, n_entries * sizeof(DevEntry)));
  if (n_uuid) CUDA_OK(h, up(h, h->b_uuid.d, h->b_uuid.h, n_uuid * sizeof(uint32_t)));
  if (n_rel) CUDA_OK(h, up(h, h->b_rel.d, h->b_rel.h, n_rel * sizeof(int4)));
  CUDA_OK(h, cudaMemsetAsync(h->b_dec.d, 0xff, (size_t)n_ops * sizeof(DevDecision), h->stream));
  CUDA_OK(h, cudaMemsetAsync(h->b_best.d, 0, (size_t)n_ops * sizeof(OpSync), h->stream));
  CUDA_OK(h, cudaMemsetAsync(h->b_counters.d, 0, 8 * sizeof(uint32_t), h->stream));
  KernelArgs ka;
  build_kernel_args(h, ka, n_ops, alloc_cap);
  ka.fail_cap = shard_fail_cap;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
  const int block = 256;
  CUDA_OK(h, cudaEventRecord(h->ev0, h->stream));
  for (int op = 0; op < n_ops; op++) {
    const DevPod& p = h->b_pods.h[op];
    cudaError_t le = cudaSuccess;
    if (p.op == OP_FILTER) {
      const int mode = !p.fast ? 0 : (hs_verdict_needs_order(p) ? 1 : 2);
      const int grid = std::max(1, std::min(sms * hs_sweep_blocks_per_sm(mode), (n_nodes * HS_D + block - 1) / block));
      le = (cudaError_t)hs_launch_sweep(&ka, op, mode, grid, block, h->stream);
      if (le == cudaSuccess && h->shard_world > 1) {
        int nr = g_nccl.AllReduce(&h->b_best.d[op].best, &h->b_best.d[op].best, 1, 5 /*ncclUint64*/, 2 /*ncclMax*/, h->nccl_comm, h->stream);
        if (nr != 0) { h->err = std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "?"); return HS_ERR_CUDA; }
      }
      if (le == cudaSuccess) le = (cudaError_t)hs_launch_shard_decide(&ka, op, h->stream);
      if (le == cudaSuccess) le = (cudaError_t)hs_launch_shard_explain(&ka, op, std::max(1, std::min(sms * 2, (n_nodes * HS_D + 255) / 256)), h->stream);
      h->launches += 3;
    } else {
      le = (cudaError_t)hs_launch_shard_decide(&ka, op, h->stream);
      h->launches += 1;
    }
    if (le != cudaSuccess) { h->err = std::string("sharded launch: ") + cudaGetErrorString(le); return HS_ERR_CUDA; }
  }
  CUDA_OK(h, cudaEventRecord(h->ev1, h->stream));
  CUDA_OK(h, down(h, h->b_counters.h, h->b_counters.d, 8 * sizeof(uint32_t)));
  CUDA_OK(h, down(h, h->b_dec.h, h->b_dec.d, (size_t)n_ops * sizeof(DevDecision)));
  CUDA_OK(h, down(h, h->b_delta.h, h->b_delta.d, (size_t)n_ops * HS_D * sizeof(int4)));
  CUDA_OK(h, cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev0, h->ev1);
  h->last_ms = ms;
  const uint32_t nalloc = std::min(h->b_counters.h[0], alloc_cap);
  if (nalloc) CUDA_OK(h, down_sync(h, h->b_alloc.h, h->b_alloc.d, (size_t)nalloc * sizeof(DevAlloc)));
  const uint32_t nfail = std::min(h->b_counters.h[1], shard_fail_cap);
  h->fail_segs.clear();
  h->op_fail_row.assign(n_ops, nullptr);
  if (nfail) {
    h->fail_segs.emplace_back((size_t)nfail * n_nodes);
    CUDA_OK(h, down_sync(h, h->fail_segs.back().data(), h->b_fail.d, (size_t)nfail * n_nodes * sizeof(DevFail)));
  }
  for (int op = 0; op < n_ops; op++) {
    const DevDecision& d = h->b_dec.h[op];
    if (d.status == 1 && d.fail_slot < nfail) h->op_fail_row[op] = h->fail_segs.back().data() + (size_t)d.fail_slot * n_nodes;
  }
  return HS_OK;
}

// Fill hs_filter_result for device op `op` (the chosen NodeScore, annotations or FailedNodes).
void assemble_result(hs_handle* h, int op, hs_filter_result* out) {
  memset(out, 0, sizeof *out);
  out->node_index = -1;
  const OpMeta& m = h->metas[op];
  if (m.status != HS_OK) { out->status = m.status; return; }
  if (m.kind == OP_RELEASE) {
    // deleting a pod whose Filter (earlier in this batch) found no node: podManager never held it
    const DevPod& p = h->b_pods.h[op];
    out->status = (p.rel_node < 0 && h->b_dec.h[p.rel_ref].status != 0) ? HS_ERR_NOTFOUND : HS_OK;
    return;
  }
  if (m.kind != OP_FILTER) { out->status = HS_OK; return; }
  const DevDecision& dec = h->b_dec.h[op];
  h->r_entries.clear(); h->r_annos.clear(); h->r_failed.clear();
  if (dec.status == 0) {
    const NodeSlot& ns = h->nodes[dec.node];
    out->status = HS_FIT;
    out->node_name = ns.name.c_str();
    out->score = dec.score;
    if (h->have_cand) {
      out->node_index = h->b_candpos.h[dec.node];
    } else if (!h->has_dead) {
      out->node_index = dec.node;
    } else {
      int idx = 0;
      for (int i = 0; i < dec.node; i++) if (h->nodes[i].alive) idx++;
      out->node_index = idx;
    }
    // Common shape (one container, one NVIDIA/DCU request): the PodDevices bookkeeping below collapses to the
    // request's type key holding the rows, plus one empty list under every other base type of the node
    // (score.go:278-282), and the annotation map to four keys.  Same output as the general code, no maps.
    if (m.n_ctr == 1 && m.n_init == 0 && m.entries.size() == 1 && (m.entries[0].vendor == 0 || m.entries[0].vendor == 1) &&
        h->b_entries.h[h->b_pods.h[op].entry_off].nums > 0 && !ns.has_mig &&
        (ns.vendor_base_types.empty() || (ns.vendor_base_types.size() == 1 && ns.vendor_base_types[0] == m.entries[0].type))) {
      const EntryMeta& em = m.entries[0];
      const VendorInfo& v = h->vendors[em.vendor];
      bool placed = false;
      auto put_rows = [&]() {
        if (dec.n_alloc == 0) { h->r_entries.push_back(hs_alloc_entry{0, em.type.c_str(), nullptr, 0, 0, 0}); return; }
        for (uint32_t i = 0; i < dec.n_alloc; i++) {
          const DevAlloc& a = h->b_alloc.h[dec.alloc_off + i];
          const DevSlot& ds = ns.devs[AL_DEV(a)];
          h->r_entries.push_back(hs_alloc_entry{0, em.type.c_str(), ds.uuid.c_str(), (int32_t)ds.index, a.usedmem, (int32_t)a.usedcores});
        }
      };
      for (auto& t : ns.base_types) {
        if (!placed && em.type <= t) { put_rows(); placed = true; if (em.type == t) continue; }
        h->r_entries.push_back(hs_alloc_entry{0, t.c_str(), nullptr, 0, 0, 0});
      }
      if (!placed) put_rows();
      std::string& buf = h->r_buf;
      buf.clear();
      auto put_int = [&buf](int v) {   // "%d" without the printf machinery
        char tmp[12];
        int n = 0;
        unsigned u = v < 0 ? 0u - (unsigned)v : (unsigned)v;
        do { tmp[n++] = (char)('0' + u % 10); u /= 10; } while (u);
        if (v < 0) buf += '-';
        while (n) buf += tmp[--n];
      };
      for (uint32_t i = 0; i < dec.n_alloc; i++) {
        const DevAlloc& a = h->b_alloc.h[dec.alloc_off + i];
        buf += ns.devs[AL_DEV(a)].uuid; buf += ','; buf += em.type; buf += ',';
        put_int(a.usedmem); buf += ','; put_int((int)a.usedcores); buf += ':';
      }
      buf += ';';
      // keys in sorted order: "<vendor>-devices-allocated" < "...-to-allocate" < "hami.io/vgpu-node" < "hami.io/vgpu-time"
      h->r_annos.push_back(hs_kv{v.allocated.c_str(), buf.c_str()});
      h->r_annos.push_back(hs_kv{v.to_allocate.c_str(), buf.c_str()});
      h->r_annos.push_back(hs_kv{"hami.io/vgpu-node", ns.name.c_str()});
      h->r_annos.push_back(hs_kv{"hami.io/vgpu-time", ""});
      out->entries = h->r_entries.data(); out->n_entries = (uint32_t)h->r_entries.size();
      out->annotations = h->r_annos.data(); out->n_annotations = (uint32_t)h->r_annos.size();
      return;
    }
    // per entry: its rows
    std::vector<std::vector<hs_alloc_entry>> rows(m.entries.size());
    for (uint32_t i = 0; i < dec.n_alloc; i++) {
      const DevAlloc& a = h->b_alloc.h[dec.alloc_off + i];
      const EntryMeta& em = m.entries[AL_ENTRY(a)];
      const DevSlot& ds = ns.devs[AL_DEV(a)];
      hs_alloc_entry row{0, em.type.c_str(), ds.uuid.c_str(), (int32_t)ds.index, a.usedmem, (int32_t)a.usedcores, nullptr, 0, 0};
      if (AL_MIG(a) && ds.mig_geom > 0 && (size_t)ds.mig_geom < h->mig_geoms.size() && h->mig_geoms[ds.mig_geom].enflame) {
        // the DRS profile the enflame Fit chose (CustomInfo["profileName"] / ["drsSlice"], enflame/device.go:452-460)
        const MigGeomHost& g = h->mig_geoms[ds.mig_geom];
        if ((size_t)AL_MIG_CAND(a) < g.enf.size()) {
          row.mig_profile = g.enf[AL_MIG_CAND(a)].name.c_str();
          row.mig_start = (uint32_t)g.enf[AL_MIG_CAND(a)].size;
        }
      } else if (AL_MIG(a) && ds.mig_geom > 0 && (size_t)ds.mig_geom < h->mig_geoms.size()) {
        // CustomInfo["migProfile"] / ["migPlacement"] of the pick (nvidia/device.go:653-657)
        const MigGeomHost& g = h->mig_geoms[ds.mig_geom];
        const MigCand& mc = g.dev.cand[AL_MIG_CAND(a)];
        const uint8_t pl = mc.place[AL_MIG_PLACE(a)];
        row.mig_profile = g.profiles[mc.prof].name.c_str();
        row.mig_start = pl >> 4; row.mig_size = pl & 15u;
      }
      rows[AL_ENTRY(a)].push_back(row);
    }
    // scoreNode's PodDevices bookkeeping incl. the base-type padding (score.go:226-286,334-338)
    typedef std::vector<std::vector<hs_alloc_entry>> PSD;
    std::map<std::string, PSD> pd, initAllocs;
    for (int c = 0; c < m.n_init && c < m.n_ctr; c++) {
      int eb = m.ctr_begin[c], ee = m.ctr_begin[c + 1];
      if (eb == ee) { for (auto& t : ns.base_types) initAllocs[t].push_back({}); continue; }
      for (int e = eb; e < ee; e++) initAllocs[m.entries[e].type].push_back(rows[e]);
      for (auto& t : ns.base_types) if ((int)initAllocs[t].size() == c) initAllocs[t].push_back({});
    }
    int appIndex = 0;
    for (int c = m.n_init; c < m.n_ctr; c++) {
      int eb = m.ctr_begin[c], ee = m.ctr_begin[c + 1];
      // sums == 0 → pad only (score.go:262-272); the compiled entries keep the caller's nums
      int sums = 0;
      for (int e = eb; e < ee; e++) sums += h->b_entries.h[h->b_pods.h[op].entry_off + e].nums;
      if (sums == 0) { for (auto& t : ns.base_types) pd[t].push_back({}); appIndex++; continue; }
      for (int e = eb; e < ee; e++) pd[m.entries[e].type].push_back(rows[e]);
      for (auto& t : ns.base_types) if ((int)pd[t].size() == appIndex) pd[t].push_back({});
      appIndex++;
    }
    if (m.n_init > 0)
      for (auto& kv : initAllocs) {
        PSD merged = kv.second;
        PSD& cur = pd[kv.first];
        merged.insert(merged.end(), cur.begin(), cur.end());
        cur = merged;
      }
    // flatten (uuid == NULL marks an empty ContainerDevices at that position)
    std::map<std::string, std::vector<std::vector<const hs_alloc_entry*>>> views;
    size_t total = 0;
    for (auto& kv : pd) for (auto& c : kv.second) total += c.empty() ? 1 : c.size();
    h->r_entries.reserve(total);
    for (auto& kv : pd) {
      const char* key = h->keep(kv.first);
      auto& view = views[kv.first];
      view.resize(kv.second.size());
      for (size_t c = 0; c < kv.second.size(); c++) {
        if (kv.second[c].empty()) { h->r_entries.push_back(hs_alloc_entry{(uint32_t)c, key, nullptr, 0, 0, 0}); continue; }
        for (auto& e : kv.second[c]) {
          hs_alloc_entry x = e;
          x.container = (uint32_t)c;
          x.type_key = key;
          h->r_entries.push_back(x);
          view[c].push_back(&h->r_entries.back());
        }
      }
    }
    // annotations: scheduler.go:1124-1130 + each vendor's PatchAnnotations
    std::map<std::string, std::string> annos;
    annos["hami.io/vgpu-node"] = ns.name;
    annos["hami.io/vgpu-time"] = "";
    for (auto& v : h->vendors) {
      auto it = views.find(v.word);
      if (it == views.end() || it->second.empty()) continue;
      std::string s = encode_pod_single(it->second, v.amd ? &ns : nullptr);
      if (v.word == "MLU") {  // cambricon/device.go:312-332
        annos["CAMBRICON_DSMLU_ASSIGNED"] = "false";
        for (auto& ctr : it->second)
          if (!ctr.empty()) {
            char buf[64];
            snprintf(buf, sizeof buf, "%d_%d_%d", ctr[0]->idx, ctr[0]->usedcores, ctr[0]->usedmem / 256);
            annos["CAMBRICON_DSMLU_PROFILE"] = buf;
            break;
          }
      }
      if (v.mthreads) {   // mthreads/device.go:155-177: allocated key only, card indices of the last container with devices
        annos[v.allocated] = s;
        for (auto& ctr : it->second) {
          if (ctr.empty()) continue;
          std::string value;
          for (auto* e : ctr) { if (!value.empty()) value += ","; value += std::to_string(e->idx); }
          annos["mthreads.com/gpu-index"] = value;
          annos["mthreads.com/predicate-time"] = "";
          annos["mthreads.com/predicate-node"] = ns.name;
        }
        continue;
      }
      if (v.sgpu) {
        // metax/sdevice.go:142-171: both hami keys, the pod's devices in metax's own JSON (protocol.go:118-134; compute and
        // vRam omitted when 0), the (blanked) predicate time
        annos[v.to_allocate] = s;
        annos[v.allocated] = s;
        std::string js = "[";
        for (size_t c = 0; c < it->second.size(); c++) {
          js += c ? ",[" : "[";
          for (size_t d = 0; d < it->second[c].size(); d++) {
            const hs_alloc_entry* e = it->second[c][d];
            js += d ? ",{" : "{";
            js += "\"uuid\":" + json_string(e->uuid);
            if (e->usedcores != 0) js += ",\"compute\":" + std::to_string(e->usedcores);
            if (e->usedmem != 0) js += ",\"vRam\":" + std::to_string(e->usedmem);
            js += "}";
          }
          js += "]";
        }
        annos["metax-tech.com/gpu-devices-allocated"] = js + "]";
        annos["metax-tech.com/predicate-time"] = "";
        continue;
      }
      if (v.enflame) {
        // enflame/device.go:254-300: the allocated key, gcu-assigned, the (blanked) time, the first container's index /
        // minor / slice count and assigned-containers keyed by container name
        annos[v.allocated] = s;
        annos["enflame.com/gcu-assigned"] = "false";
        annos["enflame.com/gcu-assigned-time"] = "";
        std::map<std::string, std::string> assigned;
        for (size_t c = 0; c < it->second.size(); c++) {
          if (it->second[c].empty()) continue;
          const hs_alloc_entry* chosen = it->second[c][0];
          const int slice = chosen->mig_start == 0 ? 1 : (int)chosen->mig_start;
          std::string profile_id, minor, index;
          for (const auto& d : ns.devs) {
            if (d.uuid != chosen->uuid) continue;
            minor = d.enf_minor; index = d.enf_index;
            if (d.mig_geom > 0 && (size_t)d.mig_geom < h->mig_geoms.size() && chosen->mig_profile)
              for (const auto& p : h->mig_geoms[d.mig_geom].enf) if (p.name == chosen->mig_profile) profile_id = p.id;
            break;
          }
          const std::string name = c < m.ctr_names.size() ? m.ctr_names[c] : "container-" + std::to_string(c);
          std::string o = "{\"allocated\":false,\"request\":" + std::to_string(slice);
          if (!profile_id.empty()) o += ",\"profileID\":" + json_string(profile_id.c_str());
          if (chosen->mig_profile && chosen->mig_profile[0]) o += ",\"profileName\":" + json_string(chosen->mig_profile);
          assigned[name] = o + "}";
          if (!annos.count("enflame.com/gcu-assigned-index") && !index.empty()) {
            annos["enflame.com/gcu-assigned-index"] = index;
            annos["enflame.com/gcu-assigned-id"] = index;
          }
          if (!annos.count("enflame.com/gcu-assigned-minor") && !minor.empty()) annos["enflame.com/gcu-assigned-minor"] = minor;
          if (!annos.count("enflame.com/gcu-request-size") && slice > 0) annos["enflame.com/gcu-request-size"] = std::to_string(slice);
        }
        if (!assigned.empty()) {
          std::string payload = "{";
          for (auto& kv : assigned) { if (payload.size() > 1) payload += ","; payload += json_string(kv.first.c_str()) + ":" + kv.second; }
          annos["assigned-containers"] = payload + "}";
        }
        continue;
      }
      if (v.kunlun && !v.vxpu) {     // kunlun/device.go:93-110: the allocated key; the card indices of the LAST container with devices
        annos[v.allocated] = s;
        for (auto& ctr : it->second) {
          std::string value;
          for (auto* e : ctr) { if (!value.empty()) value += ","; value += std::to_string(e->idx); }
          if (!value.empty()) annos["BAIDU_COM_DEVICE_IDX"] = value;
        }
        continue;
      }
      if (v.neuron) {     // awsneuron/device.go:200-243: the allocated key; the AWS_NEURON_* keys need the container's resource
        annos[v.allocated] = s;   // limits (asked by device or by core) — the caller's own PatchAnnotations adds them from the entries
        continue;
      }
      annos[v.to_allocate] = s;
      annos[v.allocated] = s;
      if (v.word == "NVIDIA") {   // EncodeMigAllocations, nvidia/device.go:511-513, nvidia/mig_allocations.go:48-76
        std::string mig;
        for (size_t c = 0; c < it->second.size(); c++)
          for (size_t di = 0; di < it->second[c].size(); di++) {
            const hs_alloc_entry* e = it->second[c][di];
            if (!e->mig_profile || !e->mig_profile[0] || e->mig_size == 0) continue;
            mig += mig.empty() ? "[" : ",";
            mig += "{\"containerIndex\":" + std::to_string(c) + ",\"deviceIndex\":" + std::to_string(di) + ",\"gpuUUID\":" +
                   json_string(e->uuid) + ",\"profile\":" + json_string(e->mig_profile) + ",\"placement\":{\"start\":" +
                   std::to_string(e->mig_start) + ",\"size\":" + std::to_string(e->mig_size) + "}}";
          }
        if (!mig.empty()) annos["hami.io/vgpu-mig-allocations"] = mig + "]";
      }
      if (v.ascend) {  // ascend/device.go:221-253
        annos["predicate-time"] = "";
        std::string json;
        bool anyrow = false;
        for (auto& ctr : it->second)
          for (auto* e : ctr) {
            std::string o = "{";
            bool first = true;
            auto add = [&](const std::string& f) { if (!first) o += ","; o += f; first = false; };
            if (e->uuid && e->uuid[0]) add(std::string("\"UUID\":\"") + e->uuid + "\"");
            if (m.vnpu_mode != "hami-core") {
              std::string temp;
              for (auto& t : v.templates) if ((int64_t)e->usedmem <= t.second) { temp = t.first; break; }
              if (!temp.empty()) add("\"temp\":\"" + temp + "\"");
            }
            if (e->usedmem != 0) add("\"memory\":" + std::to_string((long long)e->usedmem));
            if (e->usedcores != 0) add("\"core\":" + std::to_string(e->usedcores));
            o += "}";
            json += (anyrow ? "," : "") + o;
            anyrow = true;
          }
        annos["huawei.com/" + v.word] = anyrow ? "[" + json + "]" : "null";
      }
    }
    for (auto& kv : annos) h->r_annos.push_back(hs_kv{h->keep(kv.first), h->keep(kv.second)});
  } else if (dec.status == 3) {
    out->status = HS_FIT_REMOTE;      // sharded: the chosen node lives on another rank; node_index is its global position
    out->node_index = dec.node;
    out->score = dec.score;
  } else if (dec.status == 1) {
    out->status = HS_NOFIT;
    const DevFail* fr = h->op_fail_row[op];
    auto reason_of = [&](int node) -> const char* {
      if (!fr) return "";
      return fail_string(h, fr[node]).c_str();
    };
    if (h->have_cand) {
      for (size_t i = 0; i < h->cand_names.size(); i++) {
        int s = h->cand_slot[i];
        if (s < 0) h->r_failed.push_back(hs_failed_node{h->cand_names[i].c_str(), "node unregistered"});
        else h->r_failed.push_back(hs_failed_node{h->nodes[s].name.c_str(), reason_of(s)});
      }
    } else {
      // dense walk: names and liveness come from flat arrays (NodeSlot objects are large and scattered), and a run
      // of nodes with the same 12-byte reason record reuses the string of the previous one
      const size_t nn = h->nodes.size();
      h->r_failed.resize(nn);
      size_t k = 0;
      if (!fr) {
        for (size_t s = 0; s < nn; s++) if (h->h_flags[s] & NF_ALIVE) h->r_failed[k++] = hs_failed_node{h->name_ptr[s], ""};
      } else {
        const char* last = nullptr;
        uint32_t l0 = 0xffffffffu, l1 = 0, l2 = 0, l3 = 0;
        for (size_t s = 0; s < nn; s++) {
          if (!(h->h_flags[s] & NF_ALIVE)) continue;
          const DevFail& f = fr[s];
          if (!last || f.kind_cards != l0 || f.c0 != l1 || f.c1 != l2 || f.pad != l3) {
            last = fail_string(h, f).c_str();
            l0 = f.kind_cards; l1 = f.c0; l2 = f.c1; l3 = f.pad;
          }
          h->r_failed[k++] = hs_failed_node{h->name_ptr[s], last};
        }
      }
      h->r_failed.resize(k);
    }
  } else {
    out->status = HS_ERR_CUDA;
  }
  out->entries = h->r_entries.data(); out->n_entries = (uint32_t)h->r_entries.size();
  out->annotations = h->r_annos.data(); out->n_annotations = (uint32_t)h->r_annos.size();
  out->failed = h->r_failed.data(); out->n_failed = (uint32_t)h->r_failed.size();
}

// Stage a RELEASE device op for `uid` (pods map or an earlier FILTER of this batch).
// podManager (h->pods) is NOT touched here: the pod's record goes away in absorb_results, after the device has run the
// op, so every error return in between leaves the handle as it was.  `released` remembers, for the rest of this
// compile, which known pods already have a release staged.
bool stage_release(hs_handle* h, const std::string& uid, std::unordered_map<std::string, int>& in_batch,
                   std::unordered_set<std::string>& released,
                   std::vector<DevPod>& pods, std::vector<int4>& rel, std::vector<OpMeta>& metas, int user_op) {
  DevPod p;
  memset(&p, 0, sizeof p);
  p.op = OP_RELEASE;
  p.q_idx[0] = p.q_idx[1] = -1;
  std::string rel_ns;
  auto ib = in_batch.find(uid);
  if (ib != in_batch.end()) {
    p.rel_node = -1;
    p.rel_ref = ib->second;
    rel_ns = metas[ib->second].ns;
    in_batch.erase(ib);
  } else {
    auto it = h->pods.find(uid);
    if (it != h->pods.end() && released.count(uid)) it = h->pods.end();
    if (it == h->pods.end()) {
      if (h->shard_world <= 1 || user_op < 0) return false;
      p.op = OP_NOP;                       // sharded: another rank may hold it; keep the op index aligned across ranks
      OpMeta m;
      m.kind = OP_NOP;
      m.uid = uid;
      m.user_op = user_op;
      pods.push_back(p);
      metas.push_back(m);
      return true;
    }
    p.rel_node = it->second.node;
    p.rel_off = (int32_t)rel.size();
    for (int d = 0; d < HS_D; d++) { int4 v = it->second.delta[d]; v.w = 0; rel.push_back(v); }
    if (it->second.app_class) {          // metax sGPU: the pod leaves the PodInfos of its devices (rows .w bits 2-3 = its class)
      for (int d = 0; d < HS_D; d++)
        if (it->second.n_ent[d] && ((uint32_t)h->h_dyn[inv_idx(it->second.node, d)].w & DF_SGPU)) rel[(size_t)p.rel_off + d].w = (int)it->second.app_class << 2;
      p.flags |= PF_REL_MIG;
    } else if (h->nodes[it->second.node].has_mig || !it->second.mig.empty()) {
      // MIG bookkeeping of the release (PF_REL_MIG): whether this pod was one of those that make a device fail closed,
      // and the slices still covered by the reservations of the pods that remain (pods released earlier in this
      // batch are gone; pods committed in this batch carry no reservation yet)
      uint8_t o[HS_D], b[HS_D], occ[HS_D]; int bad[HS_D];
      mig_pod_effect(h, h->nodes[it->second.node], it->second, o, b);
      released.insert(uid);
      if (it->second.mig.empty()) {      // no reservation: the pod only ever moved the fail-closed counts
        for (int d = 0; d < HS_D; d++) rel[(size_t)p.rel_off + d].w = b[d] ? 1 : 0;
      } else {
        mig_node_state(h, it->second.node, &released, occ, bad);
        for (int d = 0; d < HS_D; d++) rel[(size_t)p.rel_off + d].w = (int)((b[d] ? 1u : 0u) | 2u | ((uint32_t)occ[d] << 8));
      }
      p.flags |= PF_REL_MIG;
    }
    {   // the same rows inside the op record (RelDelta over e0): the resident owner needs no second fetch
      RelDelta rd[HS_D];
      bool fits = true;
      for (int d = 0; d < HS_D; d++) {
        const int4& v = it->second.delta[d];
        if (v.x < 0 || v.x > 65535 || v.z < -32768 || v.z > 32767) fits = false;
        rd[d].mem = v.y; rd[d].cores = (int16_t)v.z; rd[d].slots = (uint16_t)v.x;
      }
      if (!fits) { h->err = "release delta outside the op record's range"; return false; }
      memcpy(&p.e0, rd, sizeof rd);
    }
    rel_ns = it->second.ns;
    released.insert(uid);
  }
  quota_rows_for(h, rel_ns, p);
  OpMeta m;
  m.kind = OP_RELEASE;
  m.ns = rel_ns;
  m.uid = uid;
  m.user_op = user_op;
  pods.push_back(p);
  metas.push_back(m);
  return true;
}

// Apply the finished batch to the host mirrors: usage cache + podManager.
void absorb_results(hs_handle* h, int first, int last) {
  for (int op = first; op < last; op++) {
    const OpMeta& m = h->metas[op];
    const DevDecision& dec = h->b_dec.h[op];
    if (m.kind == OP_FILTER && m.status == HS_OK && dec.status == 0 && !m.dry) {
      PodRec pr;
      pr.node = dec.node;
      for (int d = 0; d < HS_D; d++) {
        pr.delta[d] = h->b_delta.h[(size_t)op * HS_D + d];
        pr.n_ent[d] = pr.delta[d].x > 0 ? 1 : 0;      // collapsed: one cached entry per device (initContainer.go:36-114)
        pr.delta[d].w = 0;
      }
      apply_host_delta(h, dec.node, pr.delta, +1);
      pr.ns = m.ns;
      quota_track(h, m.ns, dec.node, pr.delta, +1);
      if (m.sg_class == 1 || m.sg_class == 2) {            // metax sGPU: the pod joins its devices' PodInfos with its class
        pr.app_class = (uint8_t)m.sg_class;
        sgpu_apply_class(h, dec.node, pr.n_ent, m.sg_class, +1);
      }
      if (m.sg_class != 1)                                  // … and addJitteryQos leaves a QoS switch in flight (metax/sdevice.go:558-577)
        for (uint32_t i = 0; i < dec.n_alloc; i++) {
          const DevAlloc& a = h->b_alloc.h[dec.alloc_off + i];
          int4& v = h->h_dyn[inv_idx(dec.node, AL_DEV(a))];
          if (!((uint32_t)v.w & DF_SGPU) || !AL_MIG(a)) continue;
          const uint32_t expected = a.usedcores == 100 ? 0u : ((uint32_t)m.sg_qos & 7u);
          if ((uint32_t)AL_MIG_CAND(a) != expected) v.w = (int)(((uint32_t)v.w & 0x0fffffffu) | (1u << 28) | (expected << 29));
        }
      h->pods[m.uid] = pr;
      if (h->nodes[dec.node].has_mig) mig_apply_orphan(h, dec.node, pr, +1);
    } else if (m.kind == OP_RELEASE) {
      const DevPod& p = h->b_pods.h[op];
      if (p.rel_node >= 0) {
        apply_host_delta(h, p.rel_node, h->b_rel.h + p.rel_off, -1);
        quota_track(h, m.ns, p.rel_node, h->b_rel.h + p.rel_off, -1);
        auto known = h->pods.find(m.uid);  // podManager.DelPod: only now, with the device's release done
        bool full = (p.flags & PF_REL_MIG) != 0;
        if (known != h->pods.end() && known->second.app_class) sgpu_apply_class(h, p.rel_node, known->second.n_ent, known->second.app_class, -1);
        if (full && !h->nodes[p.rel_node].has_mig && (known == h->pods.end() || known->second.mig.empty())) full = false;   // class counts only
        if (known != h->pods.end()) {
          h->n_parked -= known->second.parked.size();
          if (!known->second.mig.empty()) h->n_mig_pods--;
          else if (full) { mig_apply_orphan(h, p.rel_node, known->second, -1); full = false; }   // no reservation: the counts only
          h->pods.erase(known);
        }
        if (full) mig_refresh(h, p.rel_node);
      } else {
        auto it = h->pods.find(m.uid);
        if (it != h->pods.end()) {
          apply_host_delta(h, it->second.node, it->second.delta, -1);
          quota_track(h, it->second.ns, it->second.node, it->second.delta, -1);
          h->n_parked -= it->second.parked.size();
          const int rel_node = it->second.node;
          if (it->second.app_class) sgpu_apply_class(h, rel_node, it->second.n_ent, it->second.app_class, -1);
          if (h->nodes[rel_node].has_mig) mig_apply_orphan(h, rel_node, it->second, -1);   // committed in this batch: no reservation yet
          h->pods.erase(it);
        }
      }
    }
  }
}

// DevPod::win_len for the ops [first, end): how many ops from each one on the decide kernel may take as ONE window
// (hs_window.cuh).  A window holds common-shape Filters (the register-resident fast path, no dry run) that share the node
// policy's direction, and releases of namespaces without a quota; it must contain a Filter within its first HS_WIN ops.
void mark_windows(std::vector<DevPod>& pods, size_t first) {
  const size_t end = pods.size();
  int next_len = 0, next_spread = -1, filter_in = 1 << 20;   // state of the run that starts at i+1
  for (size_t i = end; i-- > first;) {
    DevPod& p = pods[i];
    int kind = 0;                                             // 0 not eligible, 1 filter, 2 release / nop
    if (p.op == OP_FILTER) kind = (p.fast && !(p.flags & PF_DRYRUN) && !hs_is_net_scored(p)) ? 1 : 0;
    else if (p.op == OP_RELEASE) kind = (p.q_idx[0] < 0 && p.q_idx[1] < 0) ? 2 : 0;
    else kind = 2;
    if (!kind) { p.win_len = 0; next_len = 0; next_spread = -1; filter_in = 1 << 20; continue; }
    const int my_spread = kind == 1 ? ((p.flags & PF_NODE_SPREAD) ? 1 : 0) : -1;
    if (kind == 1 && next_spread >= 0 && next_spread != my_spread) { next_len = 0; filter_in = 1 << 20; }   // the direction changes: the run ends here
    next_len = std::min(255, next_len + 1);
    if (kind == 1) { next_spread = my_spread; filter_in = 0; } else filter_in = std::min(filter_in + 1, 1 << 20);
    p.win_len = filter_in < HS_WIN ? (uint8_t)next_len : (uint8_t)0;
  }
}

// Compile caller ops into device ops (implicit release of a pod that is filtered again,
// scheduler.go:1092-1094), run them, absorb the results.
static inline double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// A CUDA failure while ops are on the device leaves commits in HBM that the host mirror never absorbed (and, in the
// pipelined mode, whole chunks of them): the handle refuses further work instead of scheduling on a wrong cache.
int poison(hs_handle* h, int rc) {
  if (rc == HS_ERR_CUDA && !h->poisoned) {
    h->poisoned = true;
    h->err += " [the batch failed on the device; this handle's usage cache can no longer be trusted: destroy it and rebuild from the informers]";
  }
  return rc;
}

// One hs_run_batch in flight: the compile state (what the ops have been turned into so far) and, for long batches, the
// three-stage pipeline  compile thread -> device thread -> caller  (hs_batch_wait / hs_batch_get / hs_batch_summary):
// while the device runs chunk c the compile thread prepares chunk c+1 and absorbs the results of chunk c-1 into the
// host mirrors, and the caller can already materialise the results of every finished chunk.
struct BatchJob {
  const hs_op* ops = nullptr;
  uint32_t n_ops = 0;
  std::vector<DevPod> pods;
  std::vector<DevEntry> entries;
  std::vector<uint32_t> uuid_pool;
  std::vector<int4> rel;
  std::unordered_map<std::string, int> in_batch;
  std::unordered_set<std::string> released;   // known pods with a release staged in this batch (h->pods is untouched until absorb)
  uint32_t alloc_cap = 0;
  StageMark staged;
  Compiled c;
  // pipeline
  struct Chunk { int first = 0, last = 0; StageMark from, to; uint32_t alloc_from = 0, alloc_to = 0; };
  std::vector<Chunk> chunks;
  std::vector<uint32_t> bounds;   // caller-op boundaries of the chunks: chunk k = [bounds[k], bounds[k+1]); short chunks first,
                                  // so the device starts early, then the full chunk size
  int n_staged = 0, n_done = 0, n_absorbed = 0;
  bool compile_finished = false, active = false;
  int rc = HS_OK;
  uint32_t alloc_bound = 0;
  int max_ops = 0;
  double t_begin = 0, t_first_submit = 0, t_device_done = 0;
  std::thread t_compile, t_device;
  std::mutex mu;
  std::condition_variable cv;
};

void job_compile_range(hs_handle* h, BatchJob& j, uint32_t i0, uint32_t i1) {
  for (uint32_t i = i0; i < i1; i++) {
    const hs_pod* pod = j.ops[i].pod;
    if (!pod || !pod->uid) { h->user_status[i] = HS_ERR_INVALID; continue; }
    std::string uid = pod->uid;
    if (j.ops[i].kind == HS_OP_DELETE) {
      if (stage_release(h, uid, j.in_batch, j.released, j.pods, j.rel, h->metas, (int)i)) h->user_first_op[i] = (int)j.pods.size() - 1;
      else h->user_status[i] = HS_ERR_NOTFOUND;
      continue;
    }
    int rc = compile_pod(h, pod, j.uuid_pool, j.c);
    if (rc != HS_OK) { h->user_status[i] = rc; continue; }
    if (h->dry_run) { j.c.pod.flags |= PF_DRYRUN; j.c.meta.dry = true; }   // simulation runs before TakeAndDeletePod (scheduler.go:1088-1094)
    else stage_release(h, uid, j.in_batch, j.released, j.pods, j.rel, h->metas, -1);  // TakeAndDeletePod
    j.c.pod.entry_off = (uint32_t)j.entries.size();
    j.entries.insert(j.entries.end(), j.c.entries.begin(), j.c.entries.end());
    j.c.meta.user_op = (int)i;
    j.c.pod.alloc_base = j.alloc_cap;
    j.pods.push_back(j.c.pod);
    h->metas.push_back(j.c.meta);
    if (!h->dry_run) j.in_batch[uid] = (int)j.pods.size() - 1;
    h->user_first_op[i] = (int)j.pods.size() - 1;
    j.alloc_cap += j.c.max_alloc;
  }
}

// Copy what the last compile appended into the pinned staging buffers.  A buffer that has to grow is reallocated and then
// holds everything again — the device must be idle for that (`wait_idle` is called first when a pipeline is running).
template <class WaitIdle>
int job_stage(hs_handle* h, BatchJob& j, StageMark& mark, WaitIdle wait_idle) {
  const bool grow = j.pods.size() > h->b_pods.cap || j.entries.size() > h->b_entries.cap || j.uuid_pool.size() > h->b_uuid.cap ||
                    j.rel.size() > h->b_rel.cap;
  if (grow) wait_idle();
  const size_t cp = h->b_pods.cap, ce = h->b_entries.cap, cu = h->b_uuid.cap, cr = h->b_rel.cap;
  if (!h->b_pods.ensure(std::max<size_t>(j.pods.size(), 1)) || !h->b_entries.ensure(std::max<size_t>(j.entries.size(), 1)) ||
      !h->b_uuid.ensure(std::max<size_t>(j.uuid_pool.size(), 1)) || !h->b_rel.ensure(std::max<size_t>(j.rel.size(), 1))) {
    h->err = "out of memory for batch staging";
    return HS_ERR_CUDA;
  }
  StageMark& staged = j.staged;
  if (h->b_pods.cap != cp) { staged.pods = 0; mark.pods = 0; }
  if (h->b_entries.cap != ce) { staged.entries = 0; mark.entries = 0; }
  if (h->b_uuid.cap != cu) { staged.uuid = 0; mark.uuid = 0; }
  if (h->b_rel.cap != cr) { staged.rel = 0; mark.rel = 0; }
  if (!h->has_mig) mark_windows(j.pods, staged.pods);      // the window resolver does not carry the MIG bookkeeping of a release
  memcpy(h->b_pods.h + staged.pods, j.pods.data() + staged.pods, (j.pods.size() - staged.pods) * sizeof(DevPod));
  if (j.entries.size() > staged.entries) memcpy(h->b_entries.h + staged.entries, j.entries.data() + staged.entries, (j.entries.size() - staged.entries) * sizeof(DevEntry));
  if (j.uuid_pool.size() > staged.uuid) memcpy(h->b_uuid.h + staged.uuid, j.uuid_pool.data() + staged.uuid, (j.uuid_pool.size() - staged.uuid) * sizeof(uint32_t));
  if (j.rel.size() > staged.rel) memcpy(h->b_rel.h + staged.rel, j.rel.data() + staged.rel, (j.rel.size() - staged.rel) * sizeof(int4));
  staged.pods = j.pods.size(); staged.entries = j.entries.size(); staged.uuid = j.uuid_pool.size(); staged.rel = j.rel.size();
  return HS_OK;
}

// ---- the pipeline's two threads -------------------------------------------------------------------------------
// The device side of the pipeline.  Launches are CHAINED: up to two are queued on the stream at any time, each checking
// on the device that its predecessor reached its first op (KernelArgs::chain_in), so the GPU goes from one chunk to the
// next without waiting for the host; every launch's results (counters, decisions, deltas, allocation rows) come back
// on a second stream as soon as it ends, while its successor already runs.  The fail buffer is filled across the
// launches of a batch (the cursor travels in the chain); a launch that fills it stops early, the ones queued behind it
// do nothing, and the host drains the buffer and submits the rest again.
void job_device_loop(hs_handle* h, BatchJob* jp) {
  BatchJob& j = *jp;
  cudaSetDevice(h->device);
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  struct Flight { int k, first, last, slot; uint32_t alloc_from, alloc_to; bool last_of_chunk; };
  std::deque<Flight> inflight;
  int launch_no = 0, next_submit = 0, slot_no = 0;
  uint32_t fail_seen = 0;
  int rc = HS_OK;
  auto fail = [&](const char* what, cudaError_t e) { h->err = std::string(what) + ": " + cudaGetErrorString(e); rc = HS_ERR_CUDA; };
#define DEV_OK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail(#call, e_); return false; } } while (0)
  auto set_chain = [&](uint32_t ops_done, uint32_t cursor) -> bool {       // what the NEXT launch will read
    h->h_chain[0] = ops_done; h->h_chain[1] = cursor;
    DEV_OK(cudaMemcpyAsync(h->d_chain + 2 * (launch_no & 1), h->h_chain, 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    DEV_OK(cudaStreamSynchronize(h->stream));                                // h_chain is reused
    return true;
  };
  const double t_loop0 = now_s();
  auto submit = [&](const BatchJob::Chunk& ch, int k, int first, bool upload) -> bool {
    const int slot = slot_no++ & 3;
    const double ts0 = now_s();
    struct Done { hs_handle* h; int k; double t0, base; ~Done() { if (h->hostprof) fprintf(stderr, "[hs pipe] submit chunk %d at %.3f ms, took %.3f ms\n", k, 1e3 * (t0 - base), 1e3 * (now_s() - t0)); } } done_{h, k, ts0, t_loop0};
    // request records go up on their own stream, while the previous launch still runs (copies queued between two
    // kernels of one stream cost a copy-engine hand-over each); the launch waits for them through an event
    auto upl = [&](void* dst, const void* src, size_t bytes) { h->bytes_h2d += bytes; return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->stream3); };
    if (upload) {
      if (ch.to.pods > ch.from.pods) DEV_OK(upl(h->b_pods.d + ch.from.pods, h->b_pods.h + ch.from.pods, (ch.to.pods - ch.from.pods) * sizeof(DevPod)));
      if (ch.to.entries > ch.from.entries && range_needs_entries(h, ch.first, ch.last))
        DEV_OK(upl(h->b_entries.d + ch.from.entries, h->b_entries.h + ch.from.entries, (ch.to.entries - ch.from.entries) * sizeof(DevEntry)));
      if (ch.to.uuid > ch.from.uuid) DEV_OK(upl(h->b_uuid.d + ch.from.uuid, h->b_uuid.h + ch.from.uuid, (ch.to.uuid - ch.from.uuid) * sizeof(uint32_t)));
      if (ch.to.rel > ch.from.rel) DEV_OK(upl(h->b_rel.d + ch.from.rel, h->b_rel.h + ch.from.rel, (ch.to.rel - ch.from.rel) * sizeof(int4)));
    }
    DEV_OK(cudaMemsetAsync(h->b_counters.d + 8 * slot, 0, 8 * sizeof(uint32_t), h->stream3));
    DEV_OK(cudaEventRecord(h->cu[slot], h->stream3));
    DEV_OK(cudaStreamWaitEvent(h->stream, h->cu[slot], 0));
    KernelArgs ka;
    build_kernel_args(h, ka, ch.last, j.alloc_bound);
    ka.op_begin = first;
    ka.counters = h->b_counters.d + 8 * slot;
    ka.chain_in = h->d_chain + 2 * (launch_no & 1);
    ka.chain_out = h->d_chain + 2 * ((launch_no + 1) & 1);
    launch_no++;
    DEV_OK(cudaEventRecord(h->ck0[slot], h->stream));
    const bool fk = range_is_fast(h, first, ch.last);
    DEV_OK((cudaError_t)hs_launch_decide(&ka, fk ? h->grid_fast : h->grid_now, h->block, h->stream, fk ? 1 : 0));
    h->launches++;
    DEV_OK(cudaEventRecord(h->ck1[slot], h->stream));
    // results: behind the kernel, on the second stream
    DEV_OK(cudaStreamWaitEvent(h->stream2, h->ck1[slot], 0));
    auto back = [&](void* dst, const void* src, size_t bytes) { h->bytes_d2h += bytes; return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, h->stream2); };
    DEV_OK(back(h->b_counters.h + 8 * slot, h->b_counters.d + 8 * slot, 8 * sizeof(uint32_t)));
    DEV_OK(back(h->b_dec.h + first, h->b_dec.d + first, (size_t)(ch.last - first) * sizeof(DevDecision)));
    DEV_OK(back(h->b_delta.h + (size_t)first * HS_D, h->b_delta.d + (size_t)first * HS_D, (size_t)(ch.last - first) * HS_D * sizeof(int4)));
    if (ch.alloc_to > ch.alloc_from) DEV_OK(back(h->b_alloc.h + ch.alloc_from, h->b_alloc.d + ch.alloc_from, (size_t)(ch.alloc_to - ch.alloc_from) * sizeof(DevAlloc)));
    DEV_OK(cudaEventRecord(h->cc[slot], h->stream2));
    inflight.push_back(Flight{k, first, ch.last, slot, ch.alloc_from, ch.alloc_to, true});
    return true;
  };
  // wait for the oldest launch; returns the op it reached (== f.last unless the fail buffer filled up), -1 on error
  auto complete = [&](const Flight& f) -> int {
    cudaError_t e = cudaEventSynchronize(h->cc[f.slot]);
    if (e != cudaSuccess) { fail("cudaEventSynchronize", e); return -1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, h->ck0[f.slot], h->ck1[f.slot]);
    h->last_ms += ms;
    if (h->hostprof) fprintf(stderr, "[hs pipe] chunk %d results at %.3f ms, kernel %.3f ms (%d ops)\n", f.k, 1e3 * (now_s() - t_loop0), ms, f.last - f.first);
    const uint32_t* c = h->b_counters.h + 8 * f.slot;
    if (c[6]) { h->err = "fused multi-GPU arg-best: a peer rank never arrived"; rc = HS_ERR_CUDA; return -1; }
    const int reached = c[2] == 0 ? f.first : (int)c[2];     // a launch that found its predecessor short wrote nothing
    const uint32_t nfail = c[2] == 0 ? fail_seen : c[1];
    const DevFail* seg = nullptr;
    if (nfail > fail_seen) {
      const size_t bytes = (size_t)(nfail - fail_seen) * n_nodes * sizeof(DevFail);
      h->bytes_d2h += bytes;
      // (not on stream2: the successor's result copies are already queued there, behind ITS kernel; the launch that
      // wrote these records has finished, so any stream may read them)
      e = cudaMemcpyAsync(h->b_fail.h, h->b_fail.d + (size_t)fail_seen * n_nodes, bytes, cudaMemcpyDeviceToHost, h->stream3);
      if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream3);
      if (e != cudaSuccess) { fail("fail records", e); return -1; }
      h->fail_segs.emplace_back(h->b_fail.h, h->b_fail.h + (size_t)(nfail - fail_seen) * n_nodes);
      seg = h->fail_segs.back().data();
    }
    for (int op = f.first; op < reached; op++) {
      const DevDecision& d = h->b_dec.h[op];
      if (d.status == 1 && d.fail_slot >= fail_seen && d.fail_slot < nfail) h->op_fail_row[op] = seg + (size_t)(d.fail_slot - fail_seen) * n_nodes;
    }
    fail_seen = nfail;
    return reached;
  };

  if (!h->b_counters.ensure(32)) { h->err = "out of memory"; rc = HS_ERR_CUDA; }
  if (rc == HS_OK) set_chain(0, 0);
  j.t_first_submit = 0;
  while (rc == HS_OK) {
    // keep two launches queued whenever the compile thread has staged that far
    bool finished = false;
    while (inflight.size() < 2) {
      BatchJob::Chunk ch;
      bool have = false;
      {
        std::unique_lock<std::mutex> lk(j.mu);
        if (inflight.empty()) j.cv.wait(lk, [&] { return j.n_staged > next_submit || j.compile_finished || j.rc != HS_OK; });
        if (j.rc != HS_OK) { finished = true; break; }
        if (j.n_staged > next_submit) { ch = j.chunks[next_submit]; have = true; }
        else if (j.compile_finished && inflight.empty()) { finished = true; break; }
      }
      if (!have) break;
      if (j.t_first_submit == 0) j.t_first_submit = now_s();
      const int k = next_submit++;
      if (ch.last > ch.first) { if (!submit(ch, k, ch.first, true)) break; }
      else inflight.push_back(Flight{k, ch.first, ch.last, -1, 0, 0, true});      // a chunk without device ops
    }
    if (finished || rc != HS_OK) break;
    if (inflight.empty()) continue;
    Flight f = inflight.front();
    inflight.pop_front();
    int reached = f.slot < 0 ? f.last : complete(f);
    if (reached < 0) break;
    while (reached < f.last) {
      // the fail buffer filled up inside this launch: what was queued behind it did nothing.  Drain (done above: every
      // record up to the cap is on the host), reset the cursor, run the rest of the chunk, then submit the followers again.
      if (cudaStreamSynchronize(h->stream) != cudaSuccess || cudaStreamSynchronize(h->stream2) != cudaSuccess) { fail("drain", cudaGetLastError()); break; }
      inflight.clear();
      next_submit = f.k + 1;
      fail_seen = 0;
      if (!set_chain((uint32_t)reached, 0)) break;
      if (cudaMemsetAsync(h->b_best.d + reached, 0, sizeof(OpSync), h->stream) != cudaSuccess ||
          cudaMemsetAsync(h->b_best.d + h->sync2_base + reached, 0, sizeof(OpSync), h->stream) != cudaSuccess) { fail("resume", cudaGetLastError()); break; }
      BatchJob::Chunk ch;
      { std::lock_guard<std::mutex> lk(j.mu); ch = j.chunks[f.k]; }
      const int before = reached;
      if (!submit(ch, f.k, reached, false)) break;
      Flight g = inflight.front();
      inflight.pop_front();
      reached = complete(g);
      if (reached < 0) break;
      if (reached == before && fail_seen == 0) { h->err = "decide kernel made no progress"; rc = HS_ERR_CUDA; break; }
    }
    if (rc != HS_OK || reached < 0) break;
    {
      std::lock_guard<std::mutex> lk(j.mu);
      j.n_done = f.k + 1;
    }
    j.cv.notify_all();
  }
#undef DEV_OK
  cudaStreamSynchronize(h->stream3);      // nothing of this batch is left on the device when the loop ends, whatever ended it
  cudaStreamSynchronize(h->stream);
  cudaStreamSynchronize(h->stream2);
  if (rc != HS_OK) {
    std::lock_guard<std::mutex> lk(j.mu);
    j.rc = rc;
  }
  j.cv.notify_all();
  j.t_device_done = now_s();
}

void absorb_results(hs_handle* h, int first, int last);

void job_compile_loop(hs_handle* h, BatchJob* jp) {
  BatchJob& j = *jp;
  cudaSetDevice(h->device);
  auto absorb_ready = [&](bool block) {
    for (;;) {
      int k = -1;
      {
        std::unique_lock<std::mutex> lk(j.mu);
        if (block) j.cv.wait(lk, [&] { return j.n_done > j.n_absorbed || j.rc != HS_OK || j.n_absorbed >= j.n_staged; });
        if (j.rc == HS_OK && j.n_done > j.n_absorbed) k = j.n_absorbed;
      }
      if (k < 0) return;
      absorb_results(h, j.chunks[k].first, j.chunks[k].last);
      std::lock_guard<std::mutex> lk(j.mu);
      j.n_absorbed = k + 1;
    }
  };
  auto wait_idle = [&] {       // the device has run everything staged so far
    std::unique_lock<std::mutex> lk(j.mu);
    j.cv.wait(lk, [&] { return j.n_done >= j.n_staged || j.rc != HS_OK; });
  };
  StageMark uploaded;
  int first = 0;
  for (size_t kb = 0; kb + 1 < j.bounds.size(); kb++) {
    const uint32_t i0 = j.bounds[kb], i1 = j.bounds[kb + 1];
    const uint32_t alloc_from = j.alloc_cap;
    const double tc0 = now_s();
    job_compile_range(h, j, i0, i1);
    const double tc1 = now_s();
    int rc = job_stage(h, j, uploaded, wait_idle);
    if (h->hostprof) fprintf(stderr, "[hs pipe] compile chunk %zu: %u ops, compile %.3f ms, stage %.3f ms, staged at %.3f ms\n", kb, i1 - i0, 1e3 * (tc1 - tc0), 1e3 * (now_s() - tc1), 1e3 * (now_s() - j.t_begin));
    BatchJob::Chunk ch;
    ch.first = first; ch.last = (int)j.pods.size(); ch.from = uploaded; ch.to = j.staged;
    ch.alloc_from = alloc_from; ch.alloc_to = j.alloc_cap;
    {
      std::lock_guard<std::mutex> lk(j.mu);
      if (rc != HS_OK) j.rc = rc;
      else { j.chunks.push_back(ch); j.n_staged++; }
    }
    j.cv.notify_all();
    if (rc != HS_OK) return;
    uploaded = j.staged;
    first = ch.last;
    const double ta0 = now_s();
    absorb_ready(false);                        // whatever the device has finished in the meantime
    if (h->hostprof) fprintf(stderr, "[hs pipe] absorb after chunk %zu: %.3f ms\n", kb, 1e3 * (now_s() - ta0));
    { std::lock_guard<std::mutex> lk(j.mu); if (j.rc != HS_OK) return; }
  }
  { std::lock_guard<std::mutex> lk(j.mu); j.compile_finished = true; }
  j.cv.notify_all();
  for (;;) {                                    // drain: absorb the remaining chunks as they finish
    absorb_ready(true);
    std::lock_guard<std::mutex> lk(j.mu);
    if (j.rc != HS_OK || j.n_absorbed >= j.n_staged) return;
  }
}

// Wait for the pipeline of the current batch (if any) to finish, fold its outcome into the handle.
int job_finish(hs_handle* h) {
  BatchJob* jp = h->job.get();
  if (!jp || !jp->active) return HS_OK;
  BatchJob& j = *jp;
  if (j.t_compile.joinable()) j.t_compile.join();
  if (j.t_device.joinable()) j.t_device.join();
  j.active = false;
  const double t_end = now_s();
  if (j.rc != HS_OK) return poison(h, j.rc);
  // exposed host time: before the first chunk reached the device, and after the device finished
  h->host_s[0] += std::max(0.0, j.t_first_submit - j.t_begin);
  h->host_s[1] += std::max(0.0, j.t_device_done - j.t_first_submit);
  h->host_s[2] += std::max(0.0, t_end - j.t_device_done);
  return HS_OK;
}

// Block until the result of caller op `i` of the current batch can be read (its chunk has left the device).
int job_wait_user(hs_handle* h, uint32_t i) {
  BatchJob* jp = h->job.get();
  if (!jp || !jp->active) return HS_OK;
  BatchJob& j = *jp;
  const int k = (int)(std::upper_bound(j.bounds.begin(), j.bounds.end(), i) - j.bounds.begin()) - 1;
  std::unique_lock<std::mutex> lk(j.mu);
  j.cv.wait(lk, [&] { return j.n_done > k || j.rc != HS_OK; });
  return j.rc;
}

// Compile caller ops into device ops (implicit release of a pod that is filtered again, scheduler.go:1092-1094), run
// them, absorb the results.  `async`: a long batch returns as soon as its pipeline is running.
int run_ops(hs_handle* h, const hs_op* ops, uint32_t n_ops, bool async = false) {
  int frc = job_finish(h);
  if (frc != HS_OK) return frc;
  if (h->poisoned) { h->err = "handle is unusable after an earlier device failure (hs_destroy and rebuild it)"; return HS_ERR_CUDA; }
  const double t0 = now_s();
  h->job.reset(new BatchJob());
  BatchJob& j = *h->job;
  j.ops = ops; j.n_ops = n_ops;
  h->metas.clear();
  h->user_first_op.assign(n_ops, -1);
  h->user_status.assign(n_ops, HS_OK);
  h->quota_rows.clear();

  const bool launch_per_op = h->force_shard || (h->shard_world > 1 && (!h->xgpu_connected || h->shard_use_nccl));
  // Long batches on the plain single-GPU path are compiled in chunks: while the device runs chunk c, the host compiles
  // chunk c+1 (op indices, in-batch references and the allocation offsets are batch-wide, so the chunks are one stream
  // to the kernel).  Quota rows are gathered by the compile and uploaded once, so quota'd handles take one chunk.
  const uint32_t kChunk = h->pipe_chunk;   // user ops per chunk; HS_PIPE_CHUNK overrides (tests)
  const bool pipelined = !launch_per_op && h->shard_world <= 1 && !h->have_cand && !h->dry_run && !h->any_quota_limit && n_ops > kChunk;
  if (!pipelined) {
    job_compile_range(h, j, 0, n_ops);
    int n = (int)j.pods.size();
    if (n == 0) return HS_OK;
    StageMark up;
    int rc = job_stage(h, j, up, [] {});
    if (rc != HS_OK) return rc;
    if (!h->quota_rows.empty()) {
      if (h->shard_world > 1 || h->force_shard) { h->err = "namespace quotas are not supported on the sharded path"; return HS_ERR_UNSUPPORTED; }
      if (!h->b_quota.ensure(h->quota_rows.size() * 4)) { h->err = "out of memory for the quota table"; return HS_ERR_CUDA; }
      for (size_t i = 0; i < h->quota_rows.size(); i++) {
        const hs_handle::QuotaEnt& q = h->quota_at(h->quota_rows[i].first, h->quota_rows[i].second);
        const int32_t f = h->mem_factor[h->quota_rows[i].second & 31];
        h->b_quota.h[i * 4 + 0] = q.set[0] ? (f > 1 ? q.limit[0] * (int64_t)f : q.limit[0]) : INT64_MAX;   // quota.go:80-84
        h->b_quota.h[i * 4 + 1] = q.used[0];
        h->b_quota.h[i * 4 + 2] = q.set[1] ? q.limit[1] : INT64_MAX;
        h->b_quota.h[i * 4 + 3] = q.used[1];
      }
    }
    const double t1 = now_s();
    if (launch_per_op) rc = run_device_ops_sharded(h, n, j.entries.size(), j.uuid_pool.size(), j.rel.size(), j.alloc_cap);
    else {
      rc = device_begin(h, n, j.alloc_cap);
      if (rc == HS_OK) rc = device_range(h, 0, n, StageMark(), j.staged, j.alloc_cap, true, 0, j.alloc_cap);   // rows and deltas come back with the decisions
    }
    if (rc != HS_OK) return poison(h, rc);
    const double t2 = now_s();
    absorb_results(h, 0, n);
    const double t3 = now_s();
    h->host_s[0] += t1 - t0; h->host_s[1] += t2 - t1; h->host_s[2] += t3 - t2;
    return HS_OK;
  }

  // ---- pipelined: bounds first — nothing the device writes, the caller reads or a thread appends to may move while
  // the pipeline runs
  uint32_t alloc_bound = 0;
  size_t entry_bound = 0;
  for (uint32_t i = 0; i < n_ops; i++) {
    const hs_pod* pod = ops[i].pod;
    if (!pod || ops[i].kind == HS_OP_DELETE) continue;
    for (uint32_t ci = 0; ci < pod->n_containers; ci++) {
      entry_bound += pod->containers[ci].n_reqs;
      for (uint32_t r = 0; r < pod->containers[ci].n_reqs; r++) {
        const int32_t nums = pod->containers[ci].reqs[r].nums;
        alloc_bound += (uint32_t)(nums < 0 ? 0 : (nums > HS_D ? HS_D : nums));
      }
    }
  }
  const int max_ops = (int)(2 * (size_t)n_ops + 1);      // every Filter may carry an implicit release of the pod's earlier decision
  h->metas.reserve((size_t)max_ops);
  j.pods.reserve((size_t)max_ops);
  if (!h->b_pods.ensure((size_t)max_ops) || !h->b_entries.ensure(std::max<size_t>(entry_bound, 1)) || !h->b_rel.ensure((size_t)max_ops * HS_D)) {
    h->err = "out of memory for batch staging";
    return HS_ERR_CUDA;
  }
  int rc = device_begin(h, max_ops, alloc_bound);
  if (rc != HS_OK) return poison(h, rc);
  j.alloc_bound = alloc_bound; j.max_ops = max_ops;
  for (uint32_t at = 0, step = std::max<uint32_t>(kChunk / 4, 1); at < n_ops; at += step, step = std::min(kChunk, step * 2)) j.bounds.push_back(at);
  j.bounds.push_back(n_ops);
  j.chunks.reserve(j.bounds.size());
  j.t_begin = t0;
  j.active = true;
  j.t_device = std::thread(job_device_loop, h, &j);
  j.t_compile = std::thread(job_compile_loop, h, &j);
  if (async) return HS_OK;
  return job_finish(h);
}

// every entry point but the result readers first lets a running pipeline finish
#define HS_IDLE(h) do { int _rc = job_finish(h); if (_rc != HS_OK) return _rc; } while (0)

}  // namespace

// ================================================================================================
extern "C" {

hs_handle* hs_create(const hs_config* cfg) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("hs_create: no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                     "); this engine has no CPU fallback";
    return nullptr;
  }
  hs_handle* h = new hs_handle();
  if (cfg) {
    if (cfg->node_scheduler_policy) h->node_policy = cfg->node_scheduler_policy;
    if (cfg->gpu_scheduler_policy) h->gpu_policy = cfg->gpu_scheduler_policy;
    h->ascend_vnpu_default = cfg->ascend_hami_vnpu_core != 0;
  }
  int dev = cfg && cfg->cuda_device >= 0 ? cfg->cuda_device : -1;
  if (dev < 0) cudaGetDevice(&dev);
  h->device = dev;
  if (cudaSetDevice(dev) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess) {
    g_create_error = std::string("hs_create: ") + cudaGetErrorString(cudaGetLastError());
    delete h;
    return nullptr;
  }
  {
    bool ok = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&h->stream3, cudaStreamNonBlocking) == cudaSuccess &&
              cudaMalloc(&h->d_chain, 4 * sizeof(uint32_t)) == cudaSuccess && cudaMallocHost(&h->h_chain, 4 * sizeof(uint32_t)) == cudaSuccess;
    for (int i = 0; i < 4 && ok; i++)
      ok = cudaEventCreate(&h->ck0[i]) == cudaSuccess && cudaEventCreate(&h->ck1[i]) == cudaSuccess &&
           cudaEventCreateWithFlags(&h->cc[i], cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&h->cu[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { g_create_error = std::string("hs_create: ") + cudaGetErrorString(cudaGetLastError()); delete h; return nullptr; }
  }
  add_vendor(h, "NVIDIA", false);
  add_vendor(h, "DCU", false);
  add_vendor(h, "MLU", false);
  if (cfg)
    for (uint32_t i = 0; i < cfg->n_ascend && h->vendors.size() < 31; i++) {
      add_vendor(h, cfg->ascend[i].common_word ? cfg->ascend[i].common_word : "", true);
      VendorInfo& v = h->vendors.back();
      for (uint32_t t = 0; t < cfg->ascend[i].n_templates; t++)
        v.templates.push_back({cfg->ascend[i].templates[t].name ? cfg->ascend[i].templates[t].name : "", cfg->ascend[i].templates[t].memory});
      std::stable_sort(v.templates.begin(), v.templates.end(),
                       [](const std::pair<std::string, int64_t>& a, const std::pair<std::string, int64_t>& b) { return a.second < b.second; });
    }
  if (cfg)
    for (uint32_t i = 0; i < cfg->n_vendors && h->vendors.size() < 31; i++) {
      const hs_vendor_config& vc = cfg->vendors[i];
      if (!vc.common_word || h->vendor_by_word.count(vc.common_word)) continue;
      add_vendor(h, vc.common_word, false);
      VendorInfo& v = h->vendors.back();
      v.ascend910 = false;
      v.gate = HV_DCU;                      // iluvatar's Fit is hygon's gate list (iluvatar/device.go:267-371)
      v.fit_kind = vc.fit_kind;
      v.count_only = vc.fit_kind == HS_FIT_COUNT_ONLY || vc.fit_kind == HS_FIT_VASTAI;
      v.vastai = vc.fit_kind == HS_FIT_VASTAI;
      v.amd = vc.fit_kind == HS_FIT_AMD;
      v.mthreads = vc.fit_kind == HS_FIT_MTHREADS;
      v.whole_card = vc.fit_kind == HS_FIT_WHOLE_CARD;
      v.metax = vc.fit_kind == HS_FIT_METAX;
      v.neuron = vc.fit_kind == HS_FIT_NEURON;
      v.enflame = vc.fit_kind == HS_FIT_ENFLAME_VGCU;
      v.sgpu = vc.fit_kind == HS_FIT_METAX_SGPU;
      v.option = vc.option;
      v.kunlun = vc.fit_kind == HS_FIT_KUNLUN || vc.fit_kind == HS_FIT_KUNLUN_VXPU;
      v.vxpu = vc.fit_kind == HS_FIT_KUNLUN_VXPU;
      if (v.metax) { v.use_uuid.clear(); v.nouse_uuid.clear(); }   // checkUUID is constant true (metax/device.go:124-126)
      if (vc.use_type_key) v.use_type = vc.use_type_key;
      if (vc.nouse_type_key) v.nouse_type = vc.nouse_type_key;
      if (vc.use_uuid_key) v.use_uuid = vc.use_uuid_key;
      if (vc.nouse_uuid_key) v.nouse_uuid = vc.nouse_uuid_key;
      if (vc.to_allocate_key) v.to_allocate = vc.to_allocate_key;
      if (vc.allocated_key) v.allocated = vc.allocated_key;
    }
  h->modes = {"", "hami-core", "mig", "mps"};
  for (const auto& v : h->vendors)       // metax sGPU: the three QoS policies share the mode table (ids 4-6), so a pod's policy and a device's compare by id
    if (v.sgpu) { h->modes.push_back("best-effort"); h->modes.push_back("fixed-share"); h->modes.push_back("burst-share"); break; }
  const char* blk = getenv("HS_BLOCK");
  if (blk) { int b = atoi(blk); if (b >= 32 && b <= 512 && b % 32 == 0) h->block = b; }
  h->no_fast = getenv("HS_NO_FAST") != nullptr;
  h->no_small_direct = getenv("HS_NO_SMALL_DIRECT") != nullptr;
  h->no_resident = getenv("HS_NO_RESIDENT") != nullptr;
  h->force_shard = getenv("HS_FORCE_SHARD") != nullptr;
  h->shard_use_nccl = getenv("HS_SHARD_NCCL") != nullptr;
  h->hostprof = getenv("HS_HOSTPROF") != nullptr;
  h->windows = getenv("HS_WINDOW") != nullptr;
  h->fast_kernel = getenv("HS_NO_FAST_KERNEL") == nullptr;
  if (const char* pc = getenv("HS_PIPE_CHUNK")) { if (atoi(pc) > 0) h->pipe_chunk = (uint32_t)atoi(pc); }
  if (getenv("HS_TIMING") && h->b_timing.ensure(16)) { h->timing = true; cudaMemset(h->b_timing.d, 0, 16 * sizeof(unsigned long long)); }
  if (getenv("HS_TRACE") && h->b_trace.ensure((size_t)hs_handle::kTraceOps * hs_handle::kTraceStride)) {
    h->trace = true;
    cudaMemset(h->b_trace.d, 0, (size_t)hs_handle::kTraceOps * hs_handle::kTraceStride * sizeof(unsigned long long));
  }
  const char* fc = getenv("HS_FAIL_CAP");
  if (fc && atoi(fc) > 0) { h->fail_cap = (uint32_t)atoi(fc); h->fail_cap_fixed = true; }
  if (grow_inventory(h, cfg && cfg->reserve_nodes ? (int)cfg->reserve_nodes : HS_TILE) != HS_OK) {
    g_create_error = h->err;
    delete h;
    return nullptr;
  }
  return h;
}

void hs_destroy(hs_handle* h) {
  if (h) job_finish(h);
  if (h && h->sim) { hs_destroy(h->sim); h->sim = nullptr; }
  if (h && h->hostprof)
    fprintf(stderr, "[hs hostprof] summary: fit %.3f ms, no-fit %.3f ms, other %.3f ms over %.0f results\n", 1e3 * h->prof_s[0],
            1e3 * h->prof_s[1], 1e3 * h->prof_s[2], h->prof_s[3]);
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->d_dyn) { cudaFree(h->d_dyn); cudaFree(h->d_stat); cudaFree(h->d_flags); cudaFree(h->d_pair); }
  if (h->d_tot) cudaFree(h->d_tot);
  if (h->w_skey) { cudaFree(h->w_skey); cudaFree(h->w_rmask); cudaFree(h->w_head); cudaFree(h->w_rank); }
  if (h->w_stats) cudaFree(h->w_stats);
  h->b_pods.release(); h->b_entries.release(); h->b_uuid.release(); h->b_rel.release(); h->b_dec.release();
  h->b_alloc.release(); h->b_delta.release(); h->b_best.release(); h->b_fail.release(); h->b_counters.release();
  h->b_candpos.release(); h->b_timing.release(); h->b_trace.release(); h->b_quota.release();
  for (int p = 0; p < 8; p++)        // peer mailboxes opened with cudaIpcOpenMemHandle; our own is a plain allocation
    if (h->mbox_peer[p] && h->mbox_peer[p] != h->d_mbox) cudaIpcCloseMemHandle(h->mbox_peer[p]);
  if (h->d_mbox) cudaFree(h->d_mbox);
  if (h->d_mig_tab) cudaFree(h->d_mig_tab);
  if (h->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->nccl_comm);
  for (int i = 0; i < 4; i++) { if (h->ck0[i]) cudaEventDestroy(h->ck0[i]); if (h->ck1[i]) cudaEventDestroy(h->ck1[i]); if (h->cc[i]) cudaEventDestroy(h->cc[i]); if (h->cu[i]) cudaEventDestroy(h->cu[i]); }
  if (h->stream3) cudaStreamDestroy(h->stream3);
  if (h->d_chain) cudaFree(h->d_chain);
  if (h->h_chain) cudaFreeHost(h->h_chain);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

const char* hs_last_error(const hs_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int32_t hs_node_count(const hs_handle* h) { return h->n_alive; }

int32_t hs_node_upsert(hs_handle* h, const hs_node_info* node) {
  if (!h || !node || !node->name) return HS_ERR_INVALID;
  HS_IDLE(h);
  if (node->n_devices > HS_MAX_DEVICES_PER_NODE) { h->err = "node has more devices than HS_MAX_DEVICES_PER_NODE"; return HS_ERR_UNSUPPORTED; }
  std::vector<DevSlot> devs;
  for (uint32_t i = 0; i < node->n_devices; i++) {
    const hs_device_info& d = node->devices[i];
    DevSlot s;
    s.uuid = d.id ? d.id : ""; s.type = d.type ? d.type : ""; s.mode = d.mode ? d.mode : ""; s.vendor = d.vendor ? d.vendor : "";
    s.index = d.index; s.count = d.count; s.devmem = d.devmem; s.devcore = d.devcore; s.numa = d.numa;
    s.network_id = d.network_id; s.health = d.health != 0;
    if (s.mode == "mig") {
      if (h->shard_world > 1) { h->err = "GPUs in MIG mode are not supported on the sharded (multi-GPU) path"; return HS_ERR_UNSUPPORTED; }
      if (h->mig_geoms.empty()) { std::string e; intern_geom(h, nullptr, 0, e); }     // geometry 0: no profile at all
      s.mig_geom = intern_geom(h, d.mig_profiles, d.mig_profiles ? d.n_mig_profiles : 0, h->err);
      if (s.mig_geom < 0) return HS_ERR_UNSUPPORTED;
    }
    if (s.numa < 0 || s.numa > 255 || s.index > 255 || s.network_id > 254) { h->err = "numa/index/NetworkID out of the packed range"; return HS_ERR_UNSUPPORTED; }
    s.type_id = intern_type(h, s.type);
    if (s.type_id < 0) { h->err = "more distinct device types than HS_MAX_TYPES"; return HS_ERR_UNSUPPORTED; }
    s.mode_id = intern_mode(h, s.mode);
    if (s.mode_id < 0) { h->err = "more distinct device modes than supported"; return HS_ERR_UNSUPPORTED; }
    auto v = h->vendor_by_word.find(s.vendor);
    s.vendor_id = v == h->vendor_by_word.end() ? HV_NONE : v->second;
    if (s.vendor_id != HV_NONE && h->vendors[s.vendor_id].kunlun && (s.index > 31 || s.devcore > 32767 || s.devcore < -32768)) {
      h->err = "kunlun card outside the engine's range (Index < 32, Devcore in int16)"; return HS_ERR_UNSUPPORTED;
    }
    if (s.vendor_id != HV_NONE && h->vendors[s.vendor_id].neuron && (s.devcore > 32767 || s.devcore < 0)) {
      h->err = "AWS Neuron core mask outside the engine's int16 allocation row"; return HS_ERR_UNSUPPORTED;
    }
    if (s.vendor_id != HV_NONE && h->vendors[s.vendor_id].sgpu) {
      // metax sGPU: the registered QoS policy rides in the mode bits, the MetaLink zone in the NetworkID byte
      if (h->shard_world > 1) { h->err = "metax sGPU nodes are not supported on the sharded (multi-GPU) path"; return HS_ERR_UNSUPPORTED; }
      s.sgpu = true;
      s.mode = d.sgpu && d.sgpu->qos_policy ? d.sgpu->qos_policy : "";
      s.network_id = d.sgpu ? d.sgpu->link_zone : 0;
      if (s.network_id < 0 || s.network_id > 254) { h->err = "metax sGPU LinkZone outside 0..254"; return HS_ERR_UNSUPPORTED; }
      if (s.count > 63) { h->err = "metax sGPU device with more than 63 slots (the per-device pod class counters are 6 bits)"; return HS_ERR_UNSUPPORTED; }
      s.mode_id = intern_mode(h, s.mode);
      if (s.mode_id < 0 || s.mode_id > 6) { h->err = "more distinct device modes / QoS policies than supported"; return HS_ERR_UNSUPPORTED; }
    }
    if (d.enflame) {
      s.enf_minor = d.enflame->minor ? d.enflame->minor : ""; s.enf_index = d.enflame->index ? d.enflame->index : "";
      s.enf_max_slice = d.enflame->max_slice;
      for (uint32_t k = 0; k < d.enflame->n_profiles; k++)
        s.enf_profiles.emplace_back(d.enflame->profiles[k].key ? d.enflame->profiles[k].key : "", d.enflame->profiles[k].value ? d.enflame->profiles[k].value : "");
    }
    devs.push_back(s);
  }
  {   // enflame: ONE profile table per node, collected over its vGCU devices (collectDRSProfiles, enflame/device.go:567-619)
    std::vector<DevSlot*> enf;
    for (auto& s : devs) if (s.vendor_id != HV_NONE && h->vendors[s.vendor_id].enflame) enf.push_back(&s);
    if (!enf.empty()) {
      if (h->shard_world > 1) { h->err = "enflame vGCU nodes are not supported on the sharded (multi-GPU) path"; return HS_ERR_UNSUPPORTED; }
      if (h->mig_geoms.empty()) { std::string e; intern_geom(h, nullptr, 0, e); }
      const int g = intern_enflame(h, enf, h->err);
      if (g < 0) return HS_ERR_UNSUPPORTED;
      for (auto* s : enf) s->mig_geom = g;
    }
  }
  // canonical rule 5: vendors in sorted key order, registration order inside a vendor
  std::vector<int> perm(devs.size());
  for (size_t i = 0; i < perm.size(); i++) perm[i] = (int)i;
  std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return devs[a].vendor < devs[b].vendor; });
  std::vector<DevSlot> sorted;
  for (int i : perm) sorted.push_back(devs[i]);

  int slot;
  auto it = h->node_by_name.find(node->name);
  if (it == h->node_by_name.end()) {
    slot = (int)h->nodes.size();
    int rc = grow_inventory(h, slot + 1);
    if (rc != HS_OK) return rc;
    h->nodes.emplace_back();
    h->nodes[slot].name = node->name;
    h->name_store.emplace_back(node->name);
    h->name_ptr.push_back(h->name_store.back().c_str());
    h->node_by_name[node->name] = slot;
  } else {
    slot = it->second;
  }
  NodeSlot& ns = h->nodes[slot];
  // keep committed usage of surviving UUIDs; drop index entries of the old devices
  std::unordered_map<std::string, int4> old_usage;
  for (size_t d = 0; d < ns.devs.size(); d++) {
    old_usage[ns.devs[d].uuid] = h->h_dyn[inv_idx(slot, (int)d)];
    auto& vec = h->uuid_index[ns.devs[d].uuid];
    vec.erase(std::remove(vec.begin(), vec.end(), ((uint32_t)slot << 8) | (uint32_t)d), vec.end());
    if (vec.empty()) h->uuid_index.erase(ns.devs[d].uuid);
  }
  int4 revived[HS_D];       // usage that returns with a UUID that had left this node's annotation for a while
  for (int k = 0; k < HS_D; k++) revived[k] = make_int4(0, 0, 0, 0);
  {   // the pods of this node remember their devices by slot: follow the devices to their new slots.  A vanished UUID
      // takes its share of the usage with it — and brings it back when it is registered again: the reference joins
      // pods and node devices by UUID on every rebuild (scheduler.go:773-785), so the share is parked beside the pod.
    int new_of[HS_D];
    bool identity = true;
    for (size_t d = 0; d < ns.devs.size() && d < (size_t)HS_D; d++) {
      new_of[d] = -1;
      for (size_t k = 0; k < sorted.size(); k++) if (sorted[k].uuid == ns.devs[d].uuid) { new_of[d] = (int)k; break; }
      if (new_of[d] != (int)d) identity = false;
    }
    bool fresh[HS_D];
    bool any_fresh = false;
    for (size_t k = 0; k < sorted.size() && k < (size_t)HS_D; k++) {
      fresh[k] = true;
      for (size_t d = 0; d < ns.devs.size(); d++) if (ns.devs[d].uuid == sorted[k].uuid) fresh[k] = false;
      any_fresh = any_fresh || fresh[k];
    }
    if (!identity || (any_fresh && h->n_parked > 0))
      for (auto& kv : h->pods) {
        if (kv.second.node != slot) continue;
        int4 moved[HS_D];
        for (int k = 0; k < HS_D; k++) moved[k] = make_int4(0, 0, 0, 0);
        uint8_t ment[HS_D] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (size_t d = 0; d < ns.devs.size() && d < (size_t)HS_D; d++) {
          const int4& v = kv.second.delta[d];
          if (new_of[d] >= 0) ment[new_of[d]] = kv.second.n_ent[d];
          if (new_of[d] >= 0) moved[new_of[d]] = v;
          else if (v.x | v.y | v.z) { kv.second.parked.emplace_back(ns.devs[d].uuid, v); h->n_parked++; }
        }
        if (any_fresh && !kv.second.parked.empty())
          for (size_t k = 0; k < sorted.size() && k < (size_t)HS_D; k++) {
            if (!fresh[k]) continue;
            auto& pk = kv.second.parked;
            for (size_t i = 0; i < pk.size();) {
              if (pk[i].first != sorted[k].uuid) { i++; continue; }
              moved[k].x += pk[i].second.x; moved[k].y += pk[i].second.y; moved[k].z += pk[i].second.z;
              if (ment[k] == 0) ment[k] = 1;
              revived[k].x += pk[i].second.x; revived[k].y += pk[i].second.y; revived[k].z += pk[i].second.z;
              pk.erase(pk.begin() + (long)i);
              h->n_parked--;
            }
          }
        for (int k = 0; k < HS_D; k++) { kv.second.delta[k] = moved[k]; kv.second.n_ent[k] = ment[k]; }
      }
  }
  if (!ns.alive) { h->n_alive++; }
  ns.alive = true;
  ns.devs = sorted;
  ns.has_mig = false;
  for (const auto& d : ns.devs) if (d.mode == "mig") ns.has_mig = true;
  if (ns.has_mig) h->has_mig = true;
  ns.mig_touched = ns.mig_touched || ns.has_mig;
  ns.has_cordon = node->device_cordon != nullptr;
  ns.cordon = node->device_cordon ? node->device_cordon : "";
  ns.vnpu = node->hami_vnpu_core;
  std::set<std::string> bt;
  for (auto& d : ns.devs) bt.insert(d.type);
  ns.base_types.assign(bt.begin(), bt.end());
  ns.vendor_base_types.clear();
  for (auto& t : ns.base_types) if (h->vendor_by_word.count(t)) ns.vendor_base_types.push_back(t);
  std::set<std::string> cordoned;  // cordonedDevices, nvidia/device.go:698-713
  if (ns.has_cordon && !trimSpace(ns.cordon).empty())
    for (auto& u0 : splitComma(ns.cordon)) { std::string u = trimSpace(u0); if (!u.empty()) cordoned.insert(u); }

  bool uniform = true;
  int uval = 0; bool have = false;
  const int n = (int)ns.devs.size();
  for (int d = 0; d < HS_D; d++) {
    int4 dy = make_int4(0, 0, 0, 0), st = make_int4(0, 0, 0, 0);
    if (d < n) {
      const DevSlot& s = ns.devs[d];
      auto ou = old_usage.find(s.uuid);
      if (ou != old_usage.end()) { dy.x = ou->second.x; dy.y = ou->second.y; dy.z = ou->second.z; }
      else { dy.x = revived[d].x; dy.y = revived[d].y; dy.z = revived[d].z; }
      uint32_t df = DF_PRESENT | (s.health ? (DF_HEALTH | DF_HEALTH_REG) : 0u) | (cordoned.count(s.uuid) ? DF_CORDON : 0u) |
                    ((uint32_t)s.mode_id << 3) | ((uint32_t)(s.network_id >= 0 ? s.network_id : 255) << 8);
      if (s.sgpu) {
        df |= DF_SGPU;
        if (ou != old_usage.end()) {
          // the pods on the device stay (their classes with them); a QoS switch in flight is retired once the device
          // registers with the policy it was switching to (JitteryQosCache.Sync, metax/qos.go:37-59)
          uint32_t keep = (uint32_t)ou->second.w & 0xffff0000u;
          if (SG_JQ(keep) && SG_JQID(keep) == (uint32_t)s.mode_id) keep &= 0x0fffffffu;
          df |= keep;
        }
      }
      dy.w = (int)df;
      st.x = s.count; st.y = s.devmem; st.z = s.devcore;
      st.w = (int)((uint32_t)s.numa | ((uint32_t)s.type_id << 8) | ((uint32_t)s.index << 14) | ((uint32_t)s.vendor_id << 22) |
                   ((uint32_t)s.mig_geom << 27));
      h->uuid_index[s.uuid].push_back(((uint32_t)slot << 8) | (uint32_t)d);
    }
    h->h_dyn[inv_idx(slot, d)] = dy;
    h->h_stat[inv_idx(slot, d)] = st;
  }
  int16_t* P = h->h_pair.data() + (size_t)slot * 64;
  for (int i = 0; i < 64; i++) P[i] = 0;
  for (int i = 0; i < n; i++)
    for (int j = 0; j < n; j++) {
      int v = 0;
      if (node->pair_scores) v = node->pair_scores[(size_t)perm[i] * node->n_devices + perm[j]];
      if (v > 32767 || v < -32768) { h->err = "pair score outside int16"; return HS_ERR_UNSUPPORTED; }
      P[i * 8 + j] = (int16_t)v;
      if (i != j) { if (!have) { have = true; uval = v; } else if (v != uval) uniform = false; }
    }
  uint32_t nf = (uint32_t)n | NF_ALIVE | (uniform ? NF_PAIR_UNIFORM : 0u);
  {   // metax ScoreNode tables (metax/device.go:168-214): values for 0..31 devices ride in the node's pair tile
    int n_metax = 0;
    for (const auto& d : ns.devs) if (d.vendor_id != HV_NONE && h->vendors[d.vendor_id].metax) n_metax++;
    if (n_metax > 0 && (node->metax_losses || node->metax_scores)) {
      if (n_metax != n) { h->err = "a node with metax topology annotations must hold metax devices only (the engine's pair tile carries the tables)"; return HS_ERR_UNSUPPORTED; }
      for (int which = 0; which < 2; which++) {
        const char* text = which == 0 ? node->metax_losses : node->metax_scores;
        if (!text) continue;
        nf |= which == 0 ? NF_METAX_LOSS : NF_METAX_SCORE;
        std::vector<std::pair<int64_t, int64_t>> m;
        if (!hs_ingest::decode_int_map(text, m)) continue;          // any error: every lookup reads 0
        for (auto& kv : m) {
          if (kv.second == 0) continue;
          if (kv.first < 0) continue;                                // a device count is never negative
          if (kv.first > 31 || kv.second > 32767 || kv.second < -32768) { h->err = "metax topology table outside the engine's range (counts 0..31, int16 values)"; return HS_ERR_UNSUPPORTED; }
          P[which * 32 + (int)kv.first] = (int16_t)kv.second;
        }
      }
    }
  }
  bool vnpu = h->ascend_vnpu_default;
  if (ns.vnpu >= 0) vnpu = ns.vnpu == 1;
  if (vnpu) nf |= NF_VNPU_CORE;
  for (const auto& d : ns.devs) { if (d.type == "NVIDIA") nf |= NF_BASE_NV; if (d.type == "MLU") nf |= NF_BASE_MLU; }
  {   // kunlun ScoreNode returns 0 as soon as one device of the node is not a kunlun card (kunlun/device.go:157-160)
    bool all = !ns.devs.empty();
    for (const auto& d : ns.devs) if (d.type.find("kunlun") == std::string::npos) all = false;
    if (all) nf |= NF_ALL_KUNLUN;
  }
  h->h_flags[slot] = nf;
  {
    h->h_tot[slot] = make_int4((int)nf, host_node_score_bits(h, slot), 0, 0);
  }
  mig_refresh(h, slot);       // reservations in use and fail-closed devices, from the pods the cache holds for this node
  mark_dirty(h, slot);
  return HS_OK;
}

int32_t hs_node_remove(hs_handle* h, const char* name) {
  if (!h || !name) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(name);
  if (it == h->node_by_name.end() || !h->nodes[it->second].alive) return HS_ERR_NOTFOUND;
  int slot = it->second;
  NodeSlot& ns = h->nodes[slot];
  for (size_t d = 0; d < ns.devs.size(); d++) {
    auto& vec = h->uuid_index[ns.devs[d].uuid];
    vec.erase(std::remove(vec.begin(), vec.end(), ((uint32_t)slot << 8) | (uint32_t)d), vec.end());
    if (vec.empty()) h->uuid_index.erase(ns.devs[d].uuid);
  }
  // rmNodeDevices drops the node from the candidate set only: the pods bound to it stay in podManager, keep being
  // released against it, and count again when the node returns with the same UUIDs (getNodesUsage joins pods and nodes
  // by UUID, scheduler.go:773-785).  So the devices and their usage stay; only liveness goes.
  ns.alive = false;
  h->n_alive--;
  h->has_dead = true;
  h->h_flags[slot] &= ~NF_ALIVE;
  h->h_tot[slot].x = (int)h->h_flags[slot];
  mark_dirty(h, slot);
  return HS_OK;
}

int32_t hs_usage_apply(hs_handle* h, const char* pod_uid, const char* node, const char* uuid,
                       int32_t slots, int32_t usedmem, int32_t usedcores) {
  if (!h || !pod_uid || !node || !uuid) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(node);
  if (it == h->node_by_name.end() || !h->nodes[it->second].alive) return HS_ERR_NOTFOUND;
  int slot = it->second;
  NodeSlot& ns = h->nodes[slot];
  auto pit = h->pods.find(pod_uid);
  if (pit == h->pods.end()) {
    PodRec pr;
    pr.node = slot;
    for (int d = 0; d < HS_D; d++) pr.delta[d] = make_int4(0, 0, 0, 0);
    pit = h->pods.emplace(pod_uid, pr).first;
  }
  if (pit->second.node != slot) { h->err = "pod already holds devices on another node"; return HS_ERR_INVALID; }
  bool matched = false;
  for (size_t d = 0; d < ns.devs.size(); d++) {
    if (ns.devs[d].uuid != uuid) continue;  // scheduler.go:773-785 matches by UUID
    int4 delta[HS_D];
    for (int k = 0; k < HS_D; k++) delta[k] = make_int4(0, 0, 0, 0);
    delta[d] = make_int4(std::max(slots, 1), usedmem, usedcores, 0);
    apply_host_delta(h, slot, delta, +1);
    pit->second.delta[d].x += delta[d].x; pit->second.delta[d].y += delta[d].y; pit->second.delta[d].z += delta[d].z;
    if (pit->second.n_ent[d] < 255) pit->second.n_ent[d]++;
    matched = true;
  }
  mig_refresh(h, slot);
  mark_dirty(h, slot);
  return matched ? HS_OK : HS_ERR_NOTFOUND;
}

int32_t hs_pod_delete(hs_handle* h, const char* pod_uid) {
  if (!h || !pod_uid) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->pods.find(pod_uid);
  if (it == h->pods.end()) return HS_ERR_NOTFOUND;
  apply_host_delta(h, it->second.node, it->second.delta, -1);
  quota_track(h, it->second.ns, it->second.node, it->second.delta, -1);
  mark_dirty(h, it->second.node);
  h->n_parked -= it->second.parked.size();
  if (!it->second.mig.empty()) h->n_mig_pods--;
  if (it->second.app_class) sgpu_apply_class(h, it->second.node, it->second.n_ent, it->second.app_class, -1);
  const int del_node = it->second.node;
  h->pods.erase(it);
  mig_refresh(h, del_node);
  return HS_OK;
}

// The informer re-delivers a cached pod with its reservations decoded (pi.Pod = pod, pods.go:76-77; read by
// getNodesUsage, scheduler.go:757-842)
int32_t hs_pod_set_mig(hs_handle* h, const char* pod_uid, const hs_mig_reservation* res, uint32_t n) {
  if (!h || !pod_uid || (!res && n)) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->pods.find(pod_uid);
  if (it == h->pods.end()) return HS_ERR_NOTFOUND;
  if (!it->second.mig.empty()) h->n_mig_pods--;
  it->second.mig.clear();
  for (uint32_t i = 0; i < n; i++) {
    MigRes r;
    r.uuid = res[i].gpu_uuid ? res[i].gpu_uuid : ""; r.profile = res[i].profile ? res[i].profile : "";
    r.start = res[i].start; r.size = res[i].size;
    it->second.mig.push_back(r);
  }
  if (!it->second.mig.empty()) h->n_mig_pods++;
  h->nodes[it->second.node].mig_touched = true;
  mig_refresh(h, it->second.node);
  mark_dirty(h, it->second.node);
  return HS_OK;
}

int32_t hs_snapshot_node_mig(hs_handle* h, const char* node, uint32_t* occupied, uint8_t* health, uint32_t cap) {
  if (!h || !node) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(node);
  if (it == h->node_by_name.end()) return HS_ERR_NOTFOUND;
  const NodeSlot& ns = h->nodes[it->second];
  for (size_t d = 0; d < ns.devs.size() && d < cap; d++) {
    const uint32_t df = (uint32_t)h->h_dyn[inv_idx(it->second, (int)d)].w;
    if (occupied) occupied[d] = DF_OCC(df);
    if (health) health[d] = (df & DF_HEALTH) ? 1 : 0;
  }
  return (int32_t)ns.devs.size();
}

int32_t hs_snapshot_node_sgpu(hs_handle* h, const char* node, int32_t* online, int32_t* offline, const char** pending, uint32_t cap) {
  if (!h || !node) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(node);
  if (it == h->node_by_name.end()) return HS_ERR_NOTFOUND;
  const NodeSlot& ns = h->nodes[it->second];
  for (size_t d = 0; d < ns.devs.size() && d < cap; d++) {
    const uint32_t df = (uint32_t)h->h_dyn[inv_idx(it->second, (int)d)].w;
    const bool sg = (df & DF_SGPU) != 0;
    if (online) online[d] = sg ? (int32_t)SG_ON(df) : 0;
    if (offline) offline[d] = sg ? (int32_t)SG_OFF(df) : 0;
    if (pending) pending[d] = (sg && SG_JQ(df) && SG_JQID(df) < h->modes.size()) ? h->modes[SG_JQID(df)].c_str() : nullptr;
  }
  return (int32_t)ns.devs.size();
}

int32_t hs_debug_device_rows(hs_handle* h, const char* node, int32_t device_rows[32], int32_t mirror_rows[32]) {
  if (!h || !node || !device_rows || !mirror_rows) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(node);
  if (it == h->node_by_name.end()) return HS_ERR_NOTFOUND;
  cudaSetDevice(h->device);
  int rc = flush_inventory(h);
  if (rc != HS_OK) return rc;
  CUDA_OK(h, cudaMemcpy(device_rows, h->d_dyn + inv_idx(it->second, 0), HS_D * sizeof(int4), cudaMemcpyDeviceToHost));
  memcpy(mirror_rows, &h->h_dyn[inv_idx(it->second, 0)], HS_D * sizeof(int4));
  return HS_OK;
}

int32_t hs_rollback(hs_handle* h, const char* pod_uid) { return hs_pod_delete(h, pod_uid); }

int32_t hs_pod_get(hs_handle* h, const char* pod_uid, const char** node_name, hs_pod_usage_row* rows, uint32_t cap) {
  if (!h || !pod_uid) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->pods.find(pod_uid);
  if (it == h->pods.end()) return HS_ERR_NOTFOUND;
  const NodeSlot& ns = h->nodes[it->second.node];
  if (node_name) *node_name = ns.name.c_str();
  uint32_t n = 0;
  for (size_t d = 0; d < ns.devs.size() && d < (size_t)HS_D; d++) {
    const int4& v = it->second.delta[d];
    if (!(v.x | v.y | v.z)) continue;
    if (rows && n < cap) rows[n] = hs_pod_usage_row{ns.devs[d].uuid.c_str(), v.x, v.y, v.z};
    n++;
  }
  return (int32_t)n;
}

int32_t hs_filter(hs_handle* h, const hs_pod* pod, const char* const* node_names, uint32_t n_names,
                  hs_filter_result* out) {
  if (!h || !pod || !out) return HS_ERR_INVALID;
  HS_IDLE(h);
  cudaSetDevice(h->device);
  h->r_strs.clear();
  memset(out, 0, sizeof *out);
  out->node_index = -1;
  h->have_cand = false;
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  if (node_names) {
    if (!h->b_candpos.ensure(std::max(n_nodes, HS_TILE))) { h->err = "out of memory"; return HS_ERR_CUDA; }
    for (int i = 0; i < n_nodes; i++) h->b_candpos.h[i] = -1;
    h->cand_names.clear(); h->cand_slot.clear();
    for (uint32_t i = 0; i < n_names; i++) {
      std::string nm = node_names[i] ? node_names[i] : "";
      auto it = h->node_by_name.find(nm);
      int s = (it == h->node_by_name.end() || !h->nodes[it->second].alive) ? -1 : it->second;
      h->cand_names.push_back(nm);
      h->cand_slot.push_back(s);
      if (s >= 0) h->b_candpos.h[s] = (int32_t)i;  // a repeated name keeps its last position
    }
    h->have_cand = true;
  }
  hs_op op;
  op.kind = HS_OP_FILTER;
  op.pod = pod;
  int rc = run_ops(h, &op, 1);
  if (rc != HS_OK) { out->status = rc; h->have_cand = false; return rc; }
  if (h->user_status[0] != HS_OK) { out->status = h->user_status[0]; h->have_cand = false; return out->status; }
  assemble_result(h, h->user_first_op[0], out);
  h->have_cand = false;
  return out->status;
}

// QuotaManager.AddQuota / DelQuota for ONE managed key (quota.go:243-289); the caller maps "limits.<resource name>" to
// the vendor whose GetResourceNames() lists it (IsManagedQuota, quota.go:199-211) and to memory (0) or cores (1).
static int quota_vendor(hs_handle* h, const char* vendor) {
  if (!vendor) return -1;
  for (const auto& v : h->vendors) if (v.word == vendor) return v.id;
  return -1;
}
int32_t hs_quota_set(hs_handle* h, const char* ns, const char* vendor, int32_t which, int64_t limit) {
  if (!h || !ns || (which != 0 && which != 1)) return HS_ERR_INVALID;
  HS_IDLE(h);
  const int vid = quota_vendor(h, vendor);
  if (vid < 0) { h->err = "hs_quota_set: unknown vendor (not a managed quota)"; return HS_ERR_NOTFOUND; }
  hs_handle::QuotaEnt& q = h->quota_at(ns, vid);
  q.limit[which] = limit;
  q.set[which] = true;
  h->any_quota_limit = true;
  return HS_OK;
}
int32_t hs_quota_del(hs_handle* h, const char* ns, const char* vendor, int32_t which) {
  if (!h || !ns || (which != 0 && which != 1)) return HS_ERR_INVALID;
  HS_IDLE(h);
  const int vid = quota_vendor(h, vendor);
  if (vid < 0) return HS_ERR_NOTFOUND;
  hs_handle::QuotaEnt* q = h->quota_find(ns, vid);
  if (!q) return HS_OK;
  q->limit[which] = 0;                  // the entry and its Used stay (quota.go:281-285)
  q->set[which] = false;
  return HS_OK;
}
int32_t hs_quota_get(hs_handle* h, const char* ns, const char* vendor, int32_t which, int64_t* limit, int64_t* used, int32_t* limit_set) {
  if (!h || !ns || (which != 0 && which != 1)) return HS_ERR_INVALID;
  const int vid = quota_vendor(h, vendor);
  const hs_handle::QuotaEnt* q = vid < 0 ? nullptr : h->quota_find(ns, vid);
  if (!q) return HS_ERR_NOTFOUND;
  if (limit) *limit = q->limit[which];
  if (used) *used = q->used[which];
  if (limit_set) *limit_set = q->set[which] ? 1 : 0;
  return HS_OK;
}
int32_t hs_quota_memory_factor(hs_handle* h, const char* vendor, int32_t factor) {
  if (!h) return HS_ERR_INVALID;
  const int vid = quota_vendor(h, vendor);
  if (vid < 0 || factor < 1) return HS_ERR_INVALID;
  h->mem_factor[vid & 31] = factor;
  return HS_OK;
}

int32_t hs_filter_dry(hs_handle* h, const hs_pod* pod, const char* const* node_names, uint32_t n_names, hs_filter_result* out) {
  if (!h) return HS_ERR_INVALID;
  h->dry_run = true;
  int32_t rc = hs_filter(h, pod, node_names, n_names, out);
  h->dry_run = false;
  return rc;
}

int32_t hs_run_batch(hs_handle* h, const hs_op* ops, uint32_t n_ops) {
  if (!h || (!ops && n_ops)) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  HS_IDLE(h);
  h->have_cand = false;
  int rc = run_ops(h, ops, n_ops);
  if (rc != HS_OK) return rc;
  return (int32_t)n_ops;
}

int32_t hs_run_batch_async(hs_handle* h, const hs_op* ops, uint32_t n_ops) {
  if (!h || (!ops && n_ops)) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  HS_IDLE(h);
  h->have_cand = false;
  int rc = run_ops(h, ops, n_ops, true);
  if (rc != HS_OK) return rc;
  return (int32_t)n_ops;
}

int32_t hs_batch_wait(hs_handle* h, uint32_t i) {
  if (!h) return HS_ERR_INVALID;
  if (i == 0xffffffffu) return job_finish(h);
  if (i >= h->user_first_op.size()) return HS_ERR_INVALID;
  return job_wait_user(h, i);
}

int32_t hs_batch_get(hs_handle* h, uint32_t i, hs_filter_result* out) {
  if (!h || !out || i >= h->user_first_op.size()) return HS_ERR_INVALID;
  h->r_strs.clear();
  memset(out, 0, sizeof *out);
  out->node_index = -1;
  int wrc = job_wait_user(h, i);
  if (wrc != HS_OK) { out->status = wrc; return wrc; }
  if (h->user_status[i] != HS_OK) { out->status = h->user_status[i]; return out->status; }
  if (h->user_first_op[i] < 0) { out->status = HS_ERR_INVALID; return out->status; }
  assemble_result(h, h->user_first_op[i], out);
  return out->status;
}

static inline void fnv(uint64_t& hsh, const void* p, size_t n) {
  const unsigned char* b = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) { hsh ^= b[i]; hsh *= 1099511628211ull; }
}
static inline void fnv_str(uint64_t& hsh, const char* s) { fnv(hsh, s, strlen(s) + 1); }

int32_t hs_batch_summary(hs_handle* h, uint32_t* n_fit, uint32_t* n_nofit, uint64_t* digest) {
  if (!h) return HS_ERR_INVALID;
  const double t0 = now_s();
  uint64_t hsh = 1469598103934665603ull;
  uint32_t fit = 0, nofit = 0;
  hs_filter_result r;
  for (size_t i = 0; i < h->user_first_op.size(); i++) {
    h->r_strs.clear();
    memset(&r, 0, sizeof r);
    int wrc = job_wait_user(h, (uint32_t)i);     // a running pipeline: the results of finished chunks are read while it goes on
    if (wrc != HS_OK) return wrc;
    if (h->user_status[i] != HS_OK) r.status = h->user_status[i];
    else if (h->user_first_op[i] < 0) r.status = HS_ERR_INVALID;
    else if (h->hostprof) {
      const double a0 = now_s();
      assemble_result(h, h->user_first_op[i], &r);
      h->prof_s[r.status == HS_FIT ? 0 : (r.status == HS_NOFIT ? 1 : 2)] += now_s() - a0;
      h->prof_s[3] += 1;
    } else assemble_result(h, h->user_first_op[i], &r);
    int32_t st = r.status;
    if (!digest) {   // materialise only: count, do not hash
      if (h->user_first_op[i] >= 0 && h->metas[h->user_first_op[i]].kind == OP_FILTER) { if (st == HS_FIT) fit++; else if (st == HS_NOFIT) nofit++; }
      continue;
    }
    fnv(hsh, &st, 4);
    if (h->user_first_op[i] >= 0 && h->metas[h->user_first_op[i]].kind != OP_FILTER) continue;
    if (st == HS_FIT) {
      fit++;
      fnv_str(hsh, r.node_name);
      fnv(hsh, &r.score, 4);
      for (uint32_t a = 0; a < r.n_annotations; a++) { fnv_str(hsh, r.annotations[a].key); fnv_str(hsh, r.annotations[a].value); }
    } else if (st == HS_NOFIT) {
      nofit++;
      for (uint32_t f = 0; f < r.n_failed; f++) { fnv_str(hsh, r.failed[f].node); fnv_str(hsh, r.failed[f].reason); }
    }
  }
  if (n_fit) *n_fit = fit;
  if (n_nofit) *n_nofit = nofit;
  if (digest) *digest = hsh;
  const double t1 = now_s();
  int frc = job_finish(h);
  h->host_s[3] += t1 - t0;
  return frc;
}

/* Measurement hook (no counterpart in the reference): the inventory scan in isolation.  Compiles `pod`, then
 * runs the sweep-only kernel `iters` times over the whole inventory (no commit) and reports the average
 * device time per sweep (CUDA events on the engine's stream), the winning position and score.  This is the
 * kernel whose HBM traffic is N x 256 B per launch. */
int32_t hs_sweep_probe(hs_handle* h, const hs_pod* pod, uint32_t iters, double* ms_avg, int32_t* node_index, float* score) {
  if (!h || !pod || iters == 0) return HS_ERR_INVALID;
  HS_IDLE(h);
  cudaSetDevice(h->device);
  h->have_cand = false;
  std::vector<uint32_t> uuid_pool;
  Compiled c;
  int rc = compile_pod(h, pod, uuid_pool, c);
  h->quota_rows.clear();            // measurement / scan hook: namespace quotas are not consulted
  if (rc != HS_OK) return rc;
  c.pod.q_idx[0] = c.pod.q_idx[1] = -1;
  c.pod.entry_off = 0;
  if (!h->b_pods.ensure(1) || !h->b_entries.ensure(std::max<size_t>(c.entries.size(), 1)) || !h->b_uuid.ensure(std::max<size_t>(uuid_pool.size(), 1)) ||
      !h->b_best.ensure(1) || !h->b_counters.ensure(8)) { h->err = "out of memory"; return HS_ERR_CUDA; }
  h->b_pods.h[0] = c.pod;
  memcpy(h->b_entries.h, c.entries.data(), c.entries.size() * sizeof(DevEntry));
  if (!uuid_pool.empty()) memcpy(h->b_uuid.h, uuid_pool.data(), uuid_pool.size() * sizeof(uint32_t));
  rc = flush_inventory(h);
  if (rc != HS_OK) return rc;
  CUDA_OK(h, cudaMemcpyAsync(h->b_pods.d, h->b_pods.h, sizeof(DevPod), cudaMemcpyHostToDevice, h->stream));
  CUDA_OK(h, cudaMemcpyAsync(h->b_entries.d, h->b_entries.h, c.entries.size() * sizeof(DevEntry), cudaMemcpyHostToDevice, h->stream));
  if (!uuid_pool.empty()) CUDA_OK(h, cudaMemcpyAsync(h->b_uuid.d, h->b_uuid.h, uuid_pool.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
  CUDA_OK(h, cudaMemsetAsync(h->b_best.d, 0, sizeof(OpSync), h->stream));
  KernelArgs ka;
  build_kernel_args(h, ka, 1, 0);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
  const int block = 256;
  const int mode = !c.pod.fast ? 0 : (hs_verdict_needs_order(c.pod) ? 1 : 2);
  const int grid = std::max(1, std::min(sms * hs_sweep_blocks_per_sm(mode), (ka.n_nodes * HS_D + block - 1) / block));   // one resident wave
  cudaError_t le = (cudaError_t)hs_launch_sweep(&ka, 0, mode, grid, block, h->stream);   // warm-up
  if (le != cudaSuccess) { h->err = cudaGetErrorString(le); return HS_ERR_CUDA; }
  CUDA_OK(h, cudaEventRecord(h->ev0, h->stream));
  for (uint32_t i = 0; i < iters; i++) {
    le = (cudaError_t)hs_launch_sweep(&ka, 0, mode, grid, block, h->stream);
    if (le != cudaSuccess) { h->err = cudaGetErrorString(le); return HS_ERR_CUDA; }
  }
  h->launches += iters + 1;
  CUDA_OK(h, cudaEventRecord(h->ev1, h->stream));
  CUDA_OK(h, cudaMemcpyAsync(h->b_best.h, h->b_best.d, sizeof(OpSync), cudaMemcpyDeviceToHost, h->stream));
  CUDA_OK(h, cudaStreamSynchronize(h->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, h->ev0, h->ev1);
  if (ms_avg) *ms_avg = (double)ms / iters;
  const unsigned long long key = h->b_best.h[0].best;
  const int pos = (int)((key >> 8) & 0xffffffull) - 1;
  if (node_index) *node_index = pos;
  if (score) {
    uint32_t u = (uint32_t)(key >> 32);
    if (c.pod.flags & PF_NODE_SPREAD) u = ~u;
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
    memcpy(&f, &u, 4);
    *score = pos >= 0 ? f : 0.0f;
  }
  return pos >= 0 ? HS_FIT : HS_NOFIT;
}

/* Diagnostic: cumulative host seconds spent in request compile, device run + copies, result absorb, summary. */
int64_t hs_debug_metax_table(const char* json, int64_t index) {
  std::vector<std::pair<int64_t, int64_t>> m;
  if (!json || !hs_ingest::decode_int_map(json, m)) return 0;
  for (auto& kv : m) if (kv.first == index) return kv.second;
  return 0;
}

/* Diagnostic (HS_TRACE=1): the arrival trace of the LAST launch: [ops][stride] ns, see KernelArgs::trace. */
int32_t hs_debug_trace(hs_handle* h, uint64_t* out, uint32_t cap, uint32_t* ops, uint32_t* stride) {
  if (!h || !h->trace || !out) return HS_ERR_INVALID;
  HS_IDLE(h);
  cudaSetDevice(h->device);
  const size_t n = (size_t)hs_handle::kTraceOps * hs_handle::kTraceStride;
  if (cap < n) return HS_ERR_INVALID;
  if (cudaMemcpy(out, h->b_trace.d, n * sizeof(uint64_t), cudaMemcpyDeviceToHost) != cudaSuccess) return HS_ERR_CUDA;
  if (ops) *ops = hs_handle::kTraceOps;
  if (stride) *stride = hs_handle::kTraceStride;
  return HS_OK;
}

int32_t hs_debug_copy_bytes(hs_handle* h, uint64_t out[2]) {
  if (!h) return HS_ERR_INVALID;
  HS_IDLE(h);
  out[0] = h->bytes_h2d; out[1] = h->bytes_d2h;
  return HS_OK;
}

int32_t hs_debug_host_seconds(hs_handle* h, double out[4]) {
  if (!h) return HS_ERR_INVALID;
  HS_IDLE(h);
  for (int i = 0; i < 4; i++) out[i] = h->host_s[i];
  return HS_OK;
}

int32_t hs_shard_unique_id(void* out128) {
  std::string err;
  if (!out128 || !nccl_load(err)) { g_create_error = err; return HS_ERR_CUDA; }
  return g_nccl.GetUniqueId(out128) == 0 ? HS_OK : HS_ERR_CUDA;
}

int32_t hs_shard_init(hs_handle* h, int32_t rank, int32_t world, const void* unique_id128, int64_t pos_base) {
  if (!h || world < 1 || rank < 0 || rank >= world) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  h->shard_rank = rank; h->shard_world = world; h->pos_base = pos_base;
  if (world == 1) return HS_OK;
  if (!unique_id128) return HS_ERR_INVALID;
  if (!nccl_load(h->err)) return HS_ERR_CUDA;
  NcclId id;
  memcpy(&id, unique_id128, sizeof id);
  int nr = g_nccl.CommInitRank(&h->nccl_comm, world, id, rank);
  if (nr != 0) { h->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "?"); return HS_ERR_CUDA; }
  return HS_OK;
}

int32_t hs_shard_mailbox(hs_handle* h, void* out64) {
  if (!h || !out64) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  if (!h->d_mbox) {
    CUDA_OK(h, cudaMalloc(&h->d_mbox, 64 * sizeof(unsigned long long)));
    CUDA_OK(h, cudaMemset(h->d_mbox, 0, 64 * sizeof(unsigned long long)));
  }
  cudaIpcMemHandle_t mh;
  CUDA_OK(h, cudaIpcGetMemHandle(&mh, h->d_mbox));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
  memcpy(out64, &mh, 64);
  return HS_OK;
}

int32_t hs_shard_connect(hs_handle* h, const void* handles64, int32_t world) {
  if (!h || !handles64 || world != h->shard_world || world > 8 || !h->d_mbox) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  for (int p = 0; p < world; p++) {
    if (p == h->shard_rank) { h->mbox_peer[p] = h->d_mbox; continue; }
    cudaIpcMemHandle_t mh;
    memcpy(&mh, (const char*)handles64 + 64 * p, 64);
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, mh, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { h->err = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); return HS_ERR_CUDA; }
    h->mbox_peer[p] = static_cast<unsigned long long*>(ptr);
  }
  h->xgpu_connected = true;
  return HS_OK;
}

int32_t hs_snapshot_node(hs_handle* h, const char* node, hs_device_usage_row* rows, uint32_t cap) {
  if (!h || !node) return HS_ERR_INVALID;
  HS_IDLE(h);
  auto it = h->node_by_name.find(node);
  if (it == h->node_by_name.end() || !h->nodes[it->second].alive) return HS_ERR_NOTFOUND;
  int slot = it->second;
  // read back from HBM: the device copy is the authority the kernels commit into
  cudaSetDevice(h->device);
  if (flush_inventory(h) != HS_OK) return HS_ERR_CUDA;
  int n = (int)h->nodes[slot].devs.size();
  for (int d = 0; d < n && (uint32_t)d < cap; d++) {
    int4 v;
    if (cudaMemcpy(&v, h->d_dyn + inv_idx(slot, d), sizeof v, cudaMemcpyDeviceToHost) != cudaSuccess) return HS_ERR_CUDA;
    rows[d].used = v.x; rows[d].usedmem = v.y; rows[d].usedcores = v.z;
  }
  return n;
}

extern "C" long long hs_launch_divcheck(unsigned long long n, unsigned long long seed);
/* Self-check: the branch-free float32 divide of the fast path vs the IEEE divide on n operand pairs
 * (the first 2^24 are an exhaustive small grid, the rest pseudo-random int32 pairs). Returns mismatches. */
long long hs_debug_divcheck(unsigned long long n, unsigned long long seed) { return hs_launch_divcheck(n, seed); }

/* Diagnostic (HS_TIMING=1): cumulative SM clocks block 0 spent per phase of the decide loop:
 * 0 stage, 1 release ops, 2 sweep, 3 reduce, 4 barrier, 5 winner read, 6 commit/explain, 7 whole windows;
 * inside the windows (resolver CTA): 8 op staging + sweep, 9 waiting for the arrivals, 10 pre-tracking, 11 the ops in order. */
int32_t hs_debug_timing(hs_handle* h, unsigned long long out[8]) {
  if (!h || !h->timing) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  if (cudaMemcpy(h->b_timing.h, h->b_timing.d, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return HS_ERR_CUDA;
  for (int i = 0; i < 8; i++) out[i] = h->b_timing.h[i];
  return HS_OK;
}
/* … all 16 phase clocks, and the window counters: [0] windows, [1] ops decided in windows, [2] stops on a no-fit op,
 * [3] stops on a full tracker, [4] winners that had to be loaded in the middle of a window. */
int32_t hs_debug_window_stats(hs_handle* h, unsigned long long timing16[16], uint32_t stats8[8]) {
  if (!h) return HS_ERR_INVALID;
  cudaSetDevice(h->device);
  for (int i = 0; i < 16; i++) if (timing16) timing16[i] = 0;
  for (int i = 0; i < 8; i++) if (stats8) stats8[i] = 0;
  if (timing16 && h->timing && cudaMemcpy(timing16, h->b_timing.d, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return HS_ERR_CUDA;
  if (stats8 && h->w_stats && cudaMemcpy(stats8, h->w_stats, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost) != cudaSuccess) return HS_ERR_CUDA;
  return HS_OK;
}

/* Diagnostic: when each FILTER decision of the last batch became final on the device — %globaltimer nanoseconds,
 * one per caller op (0 for ops that are not Filters or never reached the device). */
int32_t hs_debug_decision_ns(hs_handle* h, uint64_t* out, uint32_t cap) {
  if (!h || !out) return HS_ERR_INVALID;
  HS_IDLE(h);
  uint32_t n = 0;
  for (size_t i = 0; i < h->user_first_op.size() && n < cap; i++, n++) {
    out[n] = 0;
    const int op = h->user_first_op[i];
    if (op < 0 || h->metas[op].kind != OP_FILTER || h->metas[op].status != HS_OK) continue;
    const DevDecision& d = h->b_dec.h[op];
    if (d.status == 0 || d.status == 1 || d.status == 3) out[n] = ((uint64_t)d.pad[1] << 32) | d.pad[0];
  }
  return (int32_t)n;
}

double hs_last_device_ms(const hs_handle* h) { if (h) job_finish(const_cast<hs_handle*>(h)); return h ? h->last_ms : 0.0; }
uint64_t hs_launch_count(const hs_handle* h) { return h ? h->launches : 0; }

}  // extern "C"

#include "hs_ingest_abi.inc"

// ================================================================================================
// hs_sweep_static — the one entry point of this library that allocates NOTHING.
//
// Who it is for: callers that are not allowed to call cudaMalloc at the moment they need a scan —
// a scheduler plugin running inside a process whose device memory is a fixed, pre-carved pool (every
// byte already handed out at start-up), a latency-critical admission path where an allocation could
// block on the driver's allocator lock or trigger an implicit device synchronisation, or code that has
// switched the context to a mode in which the allocator must not be entered.  Every other entry point of
// this library may grow its staging buffers on demand; this one never does.
//
// What it does: the same scan as hs_sweep_probe's kernel (one request against the whole resident
// inventory, no commit) — but its scratch memory (the request record, the per-CTA partial keys and the
// "last block" ticket) lives in the file-scope array below, which exists once per GPU that loads this
// library and stays until the process exits (HS_STATIC_SCRATCH_BYTES, ~8.5 KB).  The caller passes the
// inputs (the handle whose inventory is already in HBM, and the pod) and the output (a device buffer of
// 16 bytes: the winning packed key, then the number of CTAs that took part).
//
// Rules that follow from the shared static array:
//  * The first call on a device resolves the array's device address (cudaGetSymbolAddress with that
//    device current) — make that first call before any stream capture begins.
//  * Limit: the scan uses at most HS_STATIC_MAX_CTAS thread blocks and an inventory of at most
//    HS_STATIC_MAX_NODES nodes; larger requests are refused with HS_ERR_UNSUPPORTED and a message.
//    Only the common request shape (one request entry, no UUID use/no-use list) is accepted.
//  * Every call reuses the same array and finds the previous call's contents in it: the entry point
//    rewrites the request record and zeroes the ticket itself, and the partial keys are fully
//    overwritten before they are read.  Two calls that overlap on the GPU (different streams, or two
//    handles on one device) would race on the array — issue them one after the other.
//  * The call is synchronous: it waits for the scan before it returns (the request record is copied from a
//    host stack buffer), and it rewrites the handle's request-compile state the way hs_filter does — so it
//    is not re-entrant with any other call on the same handle.
#define HS_STATIC_MAX_CTAS 1024
#define HS_STATIC_MAX_NODES (1 << 22)
#define HS_STATIC_SCRATCH_BYTES (512 + 8 * HS_STATIC_MAX_CTAS)

static __device__ __align__(256) unsigned char g_static_scratch[HS_STATIC_SCRATCH_BYTES];
// layout: [0,160) DevPod request record | [256,260) ticket | [512, 512 + 8*MAX_CTAS) per-CTA partial keys

#include "hs_eval.cuh"
#include "hs_fast.cuh"

__global__ void __launch_bounds__(256, 4)
hs_sweep_static_kernel(const int4* __restrict__ dyn, const int4* __restrict__ stat, int4* node_tot,
                       int n_nodes, unsigned char* scratch, unsigned long long* out) {
  const DevPod& pod = *reinterpret_cast<const DevPod*>(scratch);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 256);
  unsigned long long* partial = reinterpret_cast<unsigned long long*>(scratch + 512);
  __shared__ unsigned long long s_red[8];
  __shared__ bool s_last;
  KernelArgs ka;                       // only what the evaluators read
  ka.dyn = const_cast<int4*>(dyn); ka.stat = stat; ka.node_tot = node_tot; ka.uuid_pool = nullptr;
  ka.n_nodes = n_nodes; ka.cand_pos = nullptr; ka.pos_base = 0;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, glane = lane & 7, gbase = lane & 24;
  const int gg = gtid >> 3, ngroups = (gridDim.x * blockDim.x) >> 3;
  const bool spread = pod.flags & PF_NODE_SPREAD;
  const bool lite = !hs_verdict_needs_order(pod);
  unsigned long long my_key = 0;
  DevRegs n;
  FastEval fe;
  for (int node = gg; node < n_nodes; node += ngroups) {   // n_nodes % 32 == 0: warp-uniform trip count
    const int4 nt = __ldg(node_tot + node);
    const uint32_t nf = (uint32_t)nt.x;
    const int4 a = dyn[(size_t)node * HS_D + glane];
    const int4 b = __ldg(stat + (size_t)node * HS_D + glane);
    n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
    n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
    const bool fit = (lite ? fast_fit_lite(n, glane, gbase, (uint32_t)node, nf, pod, pod.e0, ka)
                           : fast_fit(n, glane, (uint32_t)node, nf, pod, pod.e0, ka, fe)) && (nf & NF_ALIVE);
    const float sc = __int_as_float(nt.y);   // ComputeDefaultScore, maintained by the commits
    if (fit) {
      const unsigned long long key = make_key(sc, spread, (uint32_t)node);
      my_key = key > my_key ? key : my_key;
    }
  }
  unsigned long long bk = my_key;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, bk, o);
    bk = t > bk ? t : bk;
  }
  if (lane == 0) s_red[threadIdx.x >> 5] = bk;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) bk = s_red[i] > bk ? s_red[i] : bk;
    partial[blockIdx.x] = bk;
    __threadfence();
    s_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {    // the last block to finish folds the partial keys
    __threadfence();
    unsigned long long best = 0;
    for (unsigned int i = 0; i < gridDim.x; i++) {
      const unsigned long long v = reinterpret_cast<volatile unsigned long long*>(partial)[i];
      best = v > best ? v : best;
    }
    out[0] = best;
    out[1] = gridDim.x;
    *ticket = 0;                       // leave the ticket at zero for the next call as well
  }
}

static unsigned char* g_static_ptr[64] = {nullptr};   // device address of g_static_scratch, per device ordinal

extern "C" int32_t hs_sweep_static(hs_handle* h, const hs_pod* pod, void* out_device16, void* cuda_stream) {
  if (!h || !pod || !out_device16) return HS_ERR_INVALID;
  HS_IDLE(h);
  if (cudaSetDevice(h->device) != cudaSuccess) { h->err = "cudaSetDevice failed"; return HS_ERR_CUDA; }
  if (h->device < 0 || h->device >= 64) { h->err = "device ordinal out of range"; return HS_ERR_UNSUPPORTED; }
  if (!g_static_ptr[h->device]) {
    // One-time set-up on this device: in host code the array's name is only a stand-in, not a device
    // address — resolve the real address with the device current, and check the result.
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) { h->err = std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e); return HS_ERR_CUDA; }
    g_static_ptr[h->device] = static_cast<unsigned char*>(p);
  }
  unsigned char* scratch = g_static_ptr[h->device];
  const int n_nodes = (int)((h->nodes.size() + HS_TILE - 1) / HS_TILE * HS_TILE);
  if (n_nodes > HS_STATIC_MAX_NODES) {
    h->err = "hs_sweep_static: inventory exceeds HS_STATIC_MAX_NODES (4194304 nodes); use hs_sweep_probe / hs_filter";
    return HS_ERR_UNSUPPORTED;
  }
  std::vector<uint32_t> uuid_pool;     // host memory only
  Compiled c;
  h->have_cand = false;
  int rc = compile_pod(h, pod, uuid_pool, c);
  h->quota_rows.clear();            // measurement / scan hook: namespace quotas are not consulted
  if (rc != HS_OK) return rc;
  c.pod.q_idx[0] = c.pod.q_idx[1] = -1;
  if (!c.pod.fast || c.pod.e0.has_use || c.pod.e0.nouse_cnt || hs_is_net_scored(c.pod)) {
    h->err = "hs_sweep_static: only the common request shape (one request entry, no UUID use/no-use list) is supported without allocation";
    return HS_ERR_UNSUPPORTED;
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
  const int block = 256;
  const int grid = std::max(1, std::min(sms * 4, (n_nodes * HS_D + block - 1) / block));
  if (grid > HS_STATIC_MAX_CTAS) {
    h->err = "hs_sweep_static: the scan would need more thread blocks than HS_STATIC_MAX_CTAS (1024)";
    return HS_ERR_UNSUPPORTED;
  }
  cudaStream_t st = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : h->stream;
  rc = flush_inventory(h);             // copies only (pending node upserts), no allocation
  if (rc != HS_OK) return rc;
  if (st != h->stream) CUDA_OK(h, cudaStreamSynchronize(h->stream));
  // the array keeps the last call's contents: rewrite the request record, zero the ticket
  CUDA_OK(h, cudaMemcpyAsync(scratch, &c.pod, sizeof(DevPod), cudaMemcpyHostToDevice, st));
  CUDA_OK(h, cudaMemsetAsync(scratch + 256, 0, 256, st));
  hs_sweep_static_kernel<<<grid, block, 0, st>>>(h->d_dyn, h->d_stat, h->d_tot, n_nodes, scratch,
                                                 static_cast<unsigned long long*>(out_device16));
  cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess) { h->err = std::string("hs_sweep_static launch: ") + cudaGetErrorString(le); return HS_ERR_CUDA; }
  h->launches++;
  CUDA_OK(h, cudaStreamSynchronize(st));   // &c.pod is a stack buffer: the copy must have completed before we return
  return HS_OK;
}
