// hs_ingest.h — internal C++ interface of the annotation decoders (see hs_ingest.cc).
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

namespace hs_ingest {

struct MigPlacementRec { uint32_t start = 0, size = 0; };
struct MigProfileRec {     // device.MigProfile (devices.go:79-90), json: name / memoryMB / core / sliceCount / placements
  std::string name;
  int32_t memory_mb = 0, core = 0;
  uint32_t slice_count = 0;
  std::vector<MigPlacementRec> placements;
};
struct MigReservationRec { std::string gpu_uuid, profile; uint32_t start = 0, size = 0; };

struct Device {            // device.DeviceInfo, the fields the scheduler path reads (devices.go:127-144)
  std::string id, type, mode, vendor;
  uint32_t index = 0;
  int32_t count = 0, devmem = 0, devcore = 0, numa = 0;
  bool health = false;
  int32_t network_id = -1; // CustomInfo["NetworkID"] (Ascend), -1 = absent
  bool has_mig_profiles = false;
  std::vector<MigProfileRec> mig_profiles;   // DeviceInfo.MIGProfiles
  bool die_strategy = false; // CustomInfo["DeviceStrategy"] == "die" (vastai die mode)
  // metax sGPU CustomInfo (metax/protocol.go:108-112)
  bool has_sgpu = false;
  std::string sgpu_qos;
  int32_t sgpu_link_zone = 0;
  // enflame DRS CustomInfo (enflame/device.go:235-252)
  bool has_enflame = false;
  std::string enf_minor, enf_index;
  std::vector<std::pair<std::string, std::string>> enf_profiles;
  int32_t enf_max_slice = 0;
  bool has_aic = false;      // CustomInfo["AIC"] is a string: the card the die sits on (vastai/device.go:374-387)
  std::string aic;
};

struct PodEntry {          // one ContainerDevice, flattened with its PodDevices key and container index
  std::string key, uuid, type;
  uint32_t container = 0;
  bool empty = false;      // marks an empty ContainerDevices at this (key, container): no device
  int32_t usedmem = 0, usedcores = 0, slots = 0;
};

// Each decoder returns true on success; on failure `err` holds the reference's error text.
bool decode_node_csv(const std::string& s, std::vector<Device>& out, std::string& err);
bool decode_node_json(const std::string& s, std::vector<Device>& out, std::string& err);
// kind 0: cambricon GetNodeDevices (cambricon/device.go:209-232), 1: mthreads (mthreads/device.go:130-153) — the devices
// of a node derived from its capacity; false + err when the cores resource is missing or zero
bool devices_from_capacity(int kind, const std::string& node, int64_t cores, int64_t memory, std::vector<Device>& out, std::string& err);
// awsneuron GetNodeDevices (awsneuron/device.go:156-198); cores_per 0 = derive from this node; network_id = 1 on "inf" nodes else 0
bool neuron_devices_from_capacity(const std::string& node, int64_t counts, int64_t cores_total, int64_t cores_per, const char* instance_type,
                                  std::vector<Device>& out, std::string& err);
// nvidia.DecodeMigAllocations (nvidia/mig_allocations.go:78-112): the hami.io/vgpu-mig-allocations annotation; "" decodes
// to no reservation; false + err on malformed JSON, an incomplete entry, a partial runtime identity or a duplicate
// (containerIndex, deviceIndex) — the caller then treats the pod as carrying no reservation (scheduler.go:760-766)
bool decode_mig_allocations(const std::string& json, std::vector<MigReservationRec>& out, std::string& err);
// metax sGPU GetNodeDevices (metax/sdevice.go:131-140,506-526; protocol.go:50-64,96-116): metax-tech.com/node-gpu-devices
bool metax_sgpu_devices(const std::string& json, std::vector<Device>& out, std::string& err);
// enflame GetNodeDevices (enflame/device.go:198-252): the node's enflame.com/gcu-drs-capacity annotation
bool enflame_devices_from_drs(const std::string& json, const std::string& node, std::vector<Device>& out, std::string& err);
// json.Unmarshal into map[int]int as parseMetaxAnnos uses it (metax/device.go:168-183): false on any error
bool decode_int_map(const std::string& json, std::vector<std::pair<int64_t, int64_t>>& out);
// scores[i*n+j] = Scores[devs[j].id] of the pair-score record whose uuid is devs[i].id (0 where absent)
bool decode_pair_scores(const std::string& json, const std::vector<Device>& devs, std::vector<int32_t>& scores, std::string& err);
// checklist: (PodDevices key, annotation name) pairs; annos: the pod's annotations
bool decode_pod_devices(const std::vector<std::pair<std::string, std::string>>& checklist,
                        const std::vector<std::pair<std::string, std::string>>& annos, std::vector<PodEntry>& out, std::string& err);
// mode 0: CollapseInitContainerUsage, 1: AppContainersOnlyDeviceUsage.  Output: per key (sorted), devices sorted by
// UUID, container 0; a key whose collapsed list is empty yields one `empty` entry.
void collapse(int mode, uint32_t n_init, const std::vector<PodEntry>& raw, std::vector<PodEntry>& out);

}  // namespace hs_ingest
