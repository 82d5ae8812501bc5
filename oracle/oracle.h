/*
 * oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (C++17, strings and maps, no CUDA) of the HAMi scheduler's
 * Filter/Score hot path, following the reference Go sources line by line
 * (each function cites the reference file:line it restates; paths relative to
 * the reference tree root, Project-HAMi/HAMi v2.10.0).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may link, load or execute this code, and only as the
 * checker or the timed CPU baseline.  The product (k8s-vgpu-scheduler_b200/)
 * never calls into it: it has no CPU fallback.
 *
 * PARITY STATUS: pinned.  tests/test_oracle_golden.py replays the reference's
 * own known-answer tests (the JSON fixtures under tests/golden, extracted from
 * the reference's _test.go files by tools/extract_golden.py) against this code.  The Go
 * reference itself cannot be built here (no Go toolchain, no module cache).
 *
 * CANONICAL LINEARISATION — the reference is nondeterministic
 * in the places below (Go map iteration, goroutine completion order, unstable
 * sort); this oracle AND the CUDA engine both fix them the same way:
 *  1. Nodes are evaluated in NodeNames order; a node's index is its position
 *     in the candidate list.
 *  2. The winner has the best score: max, or min when the node policy is
 *     exactly "spread" (policy/node_policy.go:48-54).
 *  3. Among equal scores the HIGHEST candidate index wins (what sort.Sort +
 *     take-last gives for a stable sort of nodes arriving in order).
 *  4. One container's request types are processed in sorted map-key order
 *     (score.go:88 ranges over a Go map).
 *  5. A node's devices start in registration order, vendors in sorted vendor
 *     key order (scheduler.go:665-687 ranges over a map); <=12 devices per
 *     node is asserted so Go's sort.Sort is the stable insertion sort.
 *  6. hami.io/vgpu-time and predicate-time carry "" instead of a timestamp.
 *  7. Ascend NetworkID groups with equal candidate counts are taken in
 *     ascending NetworkID (ascend/device.go:653-663); Ascend910C cards with
 *     equal size in ascending card id idx/2 (ascend/device.go:689-702).
 *  Additionally OverrideScore sums vendor ScoreNode results in sorted type-key
 *  order (policy/node_policy.go:69 ranges over a map).
 *
 * Dynamic MIG IS restated (nvidia/device.go:602-664,811-848, nvidia/mig_topology.go, the MIG half of getNodesUsage
 * scheduler.go:757-842) and pinned on mig_topology_test.go, mig_capability_test.go and scheduler/mig_allocation_test.go.
 * Namespace ResourceQuota IS restated (QuotaManager / fitQuota, device/quota.go:62-150,
 * nvidia/device.go:666-692, cambricon/device.go:351-376) and pinned on TestFitQuota.
 */
#ifndef HAMI_ORACLE_H_
#define HAMI_ORACLE_H_

#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <string>
#include <unordered_map>
#include <vector>

namespace oracle {

// ---- dynamic MIG: pkg/device/devices.go:50-100 --------------------------------
struct MigPlacement { uint32_t Start = 0, Size = 0; };            // devices.go:50-53
struct MigProfile {                                              // devices.go:79-90
  std::string Name;
  int32_t MemoryMB = 0, Core = 0;
  uint32_t SliceCount = 0;
  std::vector<MigPlacement> Placements;
};
struct MigAllocation { std::string Profile; MigPlacement Placement; };   // devices.go:92-99 (the fields the path reads)
// nvidia.MigAllocation as the hami.io/vgpu-mig-allocations annotation carries it (nvidia/mig_allocations.go:36-45)
struct MigReservation { std::string GPUUUID, Profile; MigPlacement Placement; };

// ---- pkg/device/devices.go:106-236 ------------------------------------------
struct DeviceUsage {               // devices.go:106-125
  std::string ID;
  unsigned Index = 0;
  int32_t Used = 0, Count = 0, Usedmem = 0, Totalmem = 0, Totalcore = 0, Usedcores = 0;
  std::string Mode;
  int Numa = 0;
  std::string Type;
  bool Health = true;
  bool HasNetworkID = false;       // CustomInfo["NetworkID"] present; AWSNeuron: CustomInfo["AWSNodeType"] present …
  int NetworkID = 0;               // … and 1 when that instance type contains "inf"/"Inf" (ring), else 0
  int AwsUsage = -1;               // CustomInfo["awsusageinfo"] (awsneuron/device.go:340-355,485-488), -1 = key absent
  std::vector<MigProfile> MigProfiles;            // devices.go:116
  std::vector<MigAllocation> MigAllocationsInUse; // devices.go:117
  // enflame shared vGCU: CustomInfo["minor"], ["index"], ["profiles"] (name -> id), ["maxSlice"] (enflame/device.go:235-252)
  std::string EnfMinor, EnfIndex;
  std::vector<std::pair<std::string, std::string>> EnfProfiles;
  int EnfMaxSlice = 0;
  // metax sGPU: CustomInfo["QosPolicy"] / ["LinkZone"] (metax/protocol.go:108-112) and the application classes of the pods
  // on the device (DeviceUsage.PodInfos, metax/sdevice.go:674-690)
  bool HasSgpu = false;
  std::string QosPolicy;
  int32_t LinkZone = 0;
  int OnlinePods = 0, OfflinePods = 0;
};

struct DeviceInfo {                // devices.go:127-144
  std::string ID;
  unsigned Index = 0;
  int32_t Count = 0, Devmem = 0, Devcore = 0;
  std::string Type;
  int Numa = 0;
  std::string Mode;
  bool Health = true;
  std::string DeviceVendor;
  bool HasNetworkID = false;
  int NetworkID = 0;
  std::map<std::string, int> PairScores;  // DevicePairScore.Scores
  std::vector<MigProfile> MIGProfiles;    // devices.go:141
  std::string EnfMinor, EnfIndex;         // CustomInfo of an enflame DRS device
  std::vector<std::pair<std::string, std::string>> EnfProfiles;
  int EnfMaxSlice = 0;
  bool HasSgpu = false;                   // metax sGPU CustomInfo
  std::string QosPolicy;
  int32_t LinkZone = 0;
};

struct NodeInfo {                  // devices.go:170-174 (+ the node annotations Fit reads)
  std::string ID;
  std::map<std::string, std::vector<DeviceInfo>> Devices;  // sorted vendor key = canonical rule 5
  std::optional<std::string> DeviceCordon;                  // hami.io/device-cordon
  std::optional<std::string> HamiVnpuCore;                  // hami-vnpu-core
  std::optional<std::string> MetaxLosses, MetaxScores;      // metax-tech.com/gpu.topology.losses / .scores (metax/device.go:38-39)
};

struct ContainerDevice {           // devices.go:205-216
  int Idx = 0;
  std::string UUID, Type;
  int32_t Usedmem = 0, Usedcores = 0, Slots = 0;
  bool HasNetworkID = false;       // CustomInfo carried by the Ascend Fit (ascend/device.go:573)
  int NetworkID = 0;
  int AwsUsage = -1;               // CustomInfo["awsusageinfo"] set by the AWSNeuron Fit, -1 = absent
  bool HasMig = false;             // CustomInfo["migProfile"] / ["migPlacement"] set by the NVIDIA AddResourceUsage (nvidia/device.go:653-657)
  std::string MigProfileName;
  MigPlacement MigPlace;
  // enflame Fit's CustomInfo (enflame/device.go:452-460): profileName / profileID / minor / index / drsSlice
  bool HasDrs = false;
  std::string DrsProfileName, DrsProfileID, DrsMinor, DrsIndex;
  int DrsSlice = 0, DrsRequestMem = 0, DrsRequestCore = 0;
  // metax sGPU Fit: CustomInfo = a clone of the device's at Fit time (QosPolicy, LinkZone) + the pod's annotations
  // (metax/sdevice.go:470-482: "WorkAround" for ScoreNode)
  bool HasSgpu = false;
  std::string SgpuQos;
  int32_t SgpuLinkZone = 0;
  const std::map<std::string, std::string>* PodAnnotations = nullptr;
};

struct ContainerDeviceRequest {    // devices.go:222-228
  int32_t Nums = 0;
  std::string Type;
  int32_t Memreq = 0, MemPercentagereq = 0, Coresreq = 0;
};

using ContainerDevices = std::vector<ContainerDevice>;
using ContainerDeviceRequests = std::map<std::string, ContainerDeviceRequest>;  // sorted key = canonical rule 4
using PodSingleDevice = std::vector<ContainerDevices>;
using PodDeviceRequests = std::vector<ContainerDeviceRequests>;
using PodDevices = std::map<std::string, PodSingleDevice>;

struct Pod {                       // the parts of corev1.Pod the path reads
  std::string UID, Namespace, Name;
  std::map<std::string, std::string> Annotations;
  int NumInitContainers = 0;
  std::vector<std::string> ContainerNames;   // init containers first; "" = unnamed (enflame's assigned-containers annotation is keyed by them)
  PodDeviceRequests Reqs;          // device.Resourcereqs(pod), devices.go:639-691
};

struct DeviceScoringWeights { int64_t Slot = 1, Core = 1, Memory = 1; };  // util/device_scoring_weights.go:29-38

// ---- pkg/scheduler/policy ---------------------------------------------------
struct DeviceListsScore {          // policy/gpu_policy.go:28-32
  DeviceUsage Device;
  float Score = 0;
  void ComputeScore(const ContainerDeviceRequests& requests, const DeviceScoringWeights& w);
};

struct DeviceUsageList {           // policy/gpu_policy.go:34-40
  std::vector<DeviceListsScore> DeviceLists;
  std::string Policy;
  bool NumaBind = false;
  bool Less(int i, int j) const;
  void Sort();                     // Go sort.Sort for n <= 12: stable insertion sort
};

struct NodeUsage {                 // pkg/scheduler/nodes.go:30-34
  const NodeInfo* Info = nullptr;
  DeviceUsageList Devices;
};

struct NodeScore {                 // policy/node_policy.go:27-33
  std::string NodeID;
  PodDevices Devices;
  float Score = 0;
};

// ---- pkg/device/quota.go ------------------------------------------------------
struct Quota { int64_t Used = 0, Limit = 0; bool LimitSet = false; };   // quota.go:27-37
struct ResourceNames { std::string Memory, Core; int32_t MemoryFactor = 1; };   // devices.go ResourceNames
class QuotaManager {               // quota.go:40-43; keyed namespace -> resource name
 public:
  std::map<std::string, std::map<std::string, Quota>> Quotas;
  // FitQuota, quota.go:62-95 (the vendor lookup is done by the caller, which passes its resource names)
  bool FitQuota(const std::string& ns, int64_t memreq, const ResourceNames& names, int64_t coresreq) const;
  void AddUsage(const std::string& ns, const std::map<std::string, int64_t>& usage);     // addUsageLocked, quota.go:136-150
  void RmUsage(const std::string& ns, const std::map<std::string, int64_t>& usage);      // rmUsageLocked, quota.go:169-184
  void SetLimit(const std::string& ns, const std::string& name, int64_t limit);          // addQuotaLocked, quota.go:243-268
  void ClearLimit(const std::string& ns, const std::string& name);                       // delQuotaLocked, quota.go:271-289
};

// ---- vendor seam: pkg/device/devices.go:34-48 --------------------------------
struct FitResult {
  bool fit = false;
  ContainerDevices devs;           // tmpDevs[k.Type]
  std::string reason;
};

class Devices {
 public:
  virtual ~Devices() = default;
  virtual std::string CommonWord() const = 0;
  // `allocated` = the containers of this pod placed so far (devinput of fitInDevices), `quota` = the manager FitQuota reads
  virtual FitResult Fit(const std::vector<DeviceUsage*>& devices, const ContainerDeviceRequest& request,
                        const Pod& pod, const NodeInfo* nodeInfo, const std::string& gpuPolicyDefault,
                        const PodDevices* allocated = nullptr, const QuotaManager* quota = nullptr) const = 0;
  virtual ResourceNames GetResourceNames() const { return {CommonWord() + "/mem", CommonWord() + "/core", 1}; }
  // previous = score.SnapshotDevice(appNodeCopy.Devices): every device of the node before the app containers are placed (score.go:328)
  virtual float ScoreNode(const PodSingleDevice& podDevices, const std::string& policy, const NodeInfo* nodeInfo = nullptr,
                          const std::vector<DeviceUsage>* previous = nullptr) const { return 0; }
  // returns false where the reference returns an error (only the NVIDIA MIG branch does, nvidia/device.go:646-650)
  virtual bool AddResourceUsage(DeviceUsage& n, ContainerDevice& ctr) const;
  virtual void PatchAnnotations(const Pod& pod, std::map<std::string, std::string>& annos, const PodDevices& pd) const = 0;
  // policyNeutralScorer (policy/node_policy.go:55-65): OverrideScore weights the score by +-10000 itself
  virtual bool PolicyNeutralScore() const { return false; }
  // GetNodeDevices side effect of the metax sGPU vendor: jqCache.Sync (metax/sdevice.go:131-139, qos.go:37-59)
  virtual void SyncRegistered(const std::vector<DeviceInfo>&) const {}
  virtual bool PendingQos(const std::string& uuid, std::string& out) const { (void)uuid; (void)out; return false; }
};

struct AscendTemplate { std::string Name; int64_t Memory = 0; };
struct AscendConfig {
  std::string CommonWord;
  int64_t MemoryAllocatable = 0, MemoryCapacity = 0;
  std::vector<AscendTemplate> Templates;
};

struct Config {
  std::string NodeSchedulerPolicy = "binpack";   // cmd/scheduler/main.go:70
  std::string GPUSchedulerPolicy = "spread";     // cmd/scheduler/main.go:71
  std::vector<AscendConfig> Ascend;
  bool AscendHamiVnpuCore = false;
  // vendors with a plain Fit (hs_vendor_config): iluvatar's CommonWords, biren
  struct PlainVendor { std::string CommonWord; int FitKind = 0; std::string UseUUID, NoUseUUID, ToAllocate, Allocated, UseType, NoUseType; int Option = 0; };
  std::vector<PlainVendor> Plain;
};

struct PodInfo {                   // pkg/device/pods.go:33-40
  std::string NodeID, Namespace;
  PodDevices Devices;              // collapsed
  std::vector<MigReservation> MigAllocs;   // DecodeMigAllocations(pod.Annotations["hami.io/vgpu-mig-allocations"]), in order
  std::string AppClass;                    // pod.Annotations["metax-tech.com/sgpu-app-class"]: read through DeviceUsage.PodInfos
};

struct FilterResult {
  int status = 0;                  // HS_FIT / HS_NOFIT / HS_PASSTHROUGH / <0
  std::string error;
  int nodeIndex = -1;
  NodeScore chosen;
  std::map<std::string, std::string> annotations;
  std::vector<std::pair<std::string, std::string>> failedNodes;  // candidate order
};

struct NodeScoreResult {           // score.go:288-292
  bool fit = false;
  NodeScore score;
  std::string reason;
};

class Scheduler {
 public:
  explicit Scheduler(const Config& cfg);
  ~Scheduler();

  // nodeManager.addNode / rmNodeDevices (nodes.go:76-148)
  void AddNode(const NodeInfo& info);
  const PodInfo* GetPod(const std::string& uid) const { auto it = pods_.find(uid); return it == pods_.end() ? nullptr : &it->second; }
  bool RemoveNode(const std::string& name);
  // podManager.AddPod fed by the informer (pods.go:57-97)
  bool UsageApply(const std::string& uid, const std::string& node, const std::string& uuid,
                  int32_t slots, int32_t mem, int32_t cores);
  bool DeletePod(const std::string& uid);
  // the informer delivering a pod WITH its hami.io/vgpu-mig-allocations annotation decoded (scheduler.go:137-171 → pods.go:57-97):
  // replaces what the cache holds for the pod
  bool AddPodWithMig(const std::string& uid, const std::string& ns, const std::string& node, const PodDevices& devices,
                     const std::vector<MigReservation>& allocs);
  QuotaManager quota;              // s.quotaManager
  std::map<std::string, int64_t> CountPodDevices(const PodDevices& pd) const;   // countPodDevices, quota.go:97-117

  // Scheduler.Filter (scheduler.go:1067-1154). nodeNames == nullptr: all nodes in registration order.
  FilterResult Filter(const Pod& pod, const std::vector<std::string>* nodeNames);

  // scoreNode (score.go:294-346) on an explicit usage (what Test_calcScore drives).
  NodeScoreResult ScoreNode(const std::string& nodeID, NodeUsage& node, const Pod& pod,
                            const std::string& userNodePolicy, const DeviceScoringWeights& w) const;
  // fitInDevices (score.go:81-137)
  bool FitInDevices(NodeUsage& node, const ContainerDeviceRequests& requests, const Pod& pod,
                    PodDevices& devinput, const DeviceScoringWeights& w, std::string& reason) const;

  // buildNodeUsage + usage of every cached pod (scheduler.go:654-690,744-863)
  NodeUsage BuildNodeUsage(const NodeInfo& info, const Pod& pod) const;
  std::vector<DeviceUsage> Snapshot(const std::string& node) const;

  const Devices* Vendor(const std::string& key) const;
  const Config& config() const { return cfg_; }

  bool faithful = false;   // true: rebuild usage from the whole pod cache on every Filter (scheduler.go:744-863)
  int threads = 1;         // node fan-out (the reference runs one goroutine per node, score.go:378-406)
  size_t NodeCount() const { return nodes_.size(); }

 private:
  struct NodeEntry {
    NodeInfo info;
    std::vector<DeviceUsage> usage;  // lean mode: incremental usage, registration order
    bool alive = true;
    bool migTouched = false;         // refreshMig changed Health / reservations last time: the next call must reset them
  };
  void applyUsage(NodeEntry& ne, const PodDevices& devs, int sign, const std::string& appClass = std::string());
  void refreshMig(NodeEntry& ne);   // the MIG half of getNodesUsage (scheduler.go:757-842) for one node
  void scoreRange(const std::vector<int>& cand, size_t lo, size_t hi, const Pod& pod,
                  const std::string& userNodePolicy, const DeviceScoringWeights& w,
                  std::vector<NodeScoreResult>& out) const;
  Config cfg_;
  std::map<std::string, std::unique_ptr<Devices>> vendors_;
  std::vector<std::unique_ptr<NodeEntry>> nodes_;            // registration order (slot)
  std::unordered_map<std::string, int> nodeIndex_;
  std::unordered_map<std::string, PodInfo> pods_;           // podManager
  std::vector<std::vector<DeviceUsage>> rebuilt_;           // faithful mode: per-Filter usage rebuild
};

// ---- free functions restated from the reference (also exported for the golden tests) ----
std::string GenReason(const std::map<std::string, int>& reasons, int cards);          // common/common.go:46-53
bool CheckUUID(const std::map<std::string, std::string>& annos, const std::string& id,
               const std::string& useKey, const std::string& noUseKey);                  // devices.go:693-713
bool CheckType(const std::map<std::string, std::string>& annos, const std::string& cardType,
               const std::string& useKey, const std::string& noUseKey);                  // devices.go:719-738
bool PolicyContains(const std::string& policy, const std::string& name);                 // util/util.go:290-298
std::string GetGPUSchedulerPolicyByPod(const std::string& def, const Pod& pod);         // util/util.go:277-285
bool ParseDeviceScoringWeights(const std::string& v, DeviceScoringWeights& out, std::string& err);  // util/device_scoring_weights.go:43-102
// fitQuota (nvidia/device.go:666-692, cambricon/device.go:351-376) for the vendor `word` ("NVIDIA" / "MLU")
bool FitQuotaForDevice(const Scheduler& s, const Pod& pod, const ContainerDevices& tmpDevs, const PodDevices* allocated,
                       const std::string& word, const std::string& devUUID, int64_t memreq, int64_t coresreq);
PodDevices CollapseInitContainerUsage(int numInit, const PodDevices& raw);               // device/initContainer.go:36-114
std::string EncodePodSingleDevice(const PodSingleDevice& pd);                            // devices.go:451-484
float ComputeDefaultScore(const DeviceUsageList& devices);                               // policy/node_policy.go:106-130
// nvidia/mig_topology.go:57-79, nvidia/device.go:621-639, mig_topology.go:85-132
// kunlun/topo.go: graghSelect over list positions (fit[i] = fitFn of the device at position i, index[i] = its Index),
// calcscore on the card indices in use before / picked now
std::vector<int> KunlunGraphSelect(const std::vector<bool>& fit, const std::vector<int>& index, int count);
float KunlunCalcScore(const std::vector<int>& prev, const std::vector<int>& cur);
bool SelectMigPlacement(const std::vector<MigProfile>& profiles, const std::vector<MigPlacement>& occupied,
                        const std::string& profileName, MigPlacement& out);
bool SelectMigCandidate(const std::vector<MigProfile>& profiles, const std::vector<MigPlacement>& occupied, int32_t memory,
                        MigProfile& profile, MigPlacement& placement);
bool CanPlaceMigProfiles(const std::vector<MigProfile>& capabilities, const std::vector<MigPlacement>& occupied,
                         const std::vector<std::string>& requested);
// EncodeMigAllocations, nvidia/mig_allocations.go:48-76 ("" when there is nothing to encode)
std::string EncodeMigAllocations(const PodSingleDevice& pd);

}  // namespace oracle
#endif
