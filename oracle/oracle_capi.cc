// oracle_capi.cc — TEST INFRASTRUCTURE ONLY.  ctypes-facing wrapper over oracle.cc.
#include "oracle_capi.h"

#include <array>
#include <chrono>
#include <cstdio>
#include <algorithm>
#include <cstring>
#include <deque>
#include <stdexcept>

#include "oracle.h"

using namespace oracle;

namespace {

static const char* kReasonNames[HS_R_COUNT] = {
    "CardTypeMismatch", "CardUuidMismatch", "CardTimeSlicingExhausted", "CardComputeUnitsExhausted",
    "CardInsufficientMemory", "CardInsufficientCore", "CardNotHealth", "CardCordoned", "NumaNotFit",
    "ExclusiveDeviceAllocateConflict", "CardNotFoundCustomFilterRule", "CardMigTopologyInfeasible",
    "AllocatedCardsInsufficientRequest", "ResourceQuotaNotFit", "ModeNotFit",
    "CardUnhealthy", "CardAppClassMismatch", "CardQosPolicyMismatch"};

struct Arena {
  std::deque<std::string> strs;
  std::vector<hs_alloc_entry> entries;
  std::vector<hs_kv> annos;
  std::vector<hs_failed_node> failed;
  void clear() { strs.clear(); entries.clear(); annos.clear(); failed.clear(); }
  const char* s(const std::string& v) { strs.push_back(v); return strs.back().c_str(); }
};

thread_local Arena g_arena;

std::string str(const char* p) { return p ? std::string(p) : std::string(); }

Pod toPod(const hs_pod* p) {
  Pod pod;
  if (!p) return pod;
  pod.UID = str(p->uid);
  pod.Namespace = str(p->ns);
  pod.Name = str(p->name);
  for (uint32_t i = 0; i < p->n_annotations; i++) pod.Annotations[str(p->annotations[i].key)] = str(p->annotations[i].value);
  pod.NumInitContainers = (int)p->n_init;
  for (uint32_t c = 0; c < p->n_containers; c++) {
    ContainerDeviceRequests m;
    for (uint32_t r = 0; r < p->containers[c].n_reqs; r++) {
      const hs_container_request& q = p->containers[c].reqs[r];
      ContainerDeviceRequest k;
      k.Nums = q.nums; k.Type = str(q.type); k.Memreq = q.memreq; k.MemPercentagereq = q.mempct; k.Coresreq = q.coresreq;
      m[str(q.key)] = k;
    }
    pod.Reqs.push_back(m);
    pod.ContainerNames.push_back(p->containers[c].name ? str(p->containers[c].name) : "container-" + std::to_string(c));   // containerNameByIndex, enflame/device.go:692-705
  }
  return pod;
}

std::vector<MigProfile> toMigProfiles(const hs_mig_profile* p, uint32_t n) {
  std::vector<MigProfile> out;
  for (uint32_t i = 0; p && i < n; i++) {
    MigProfile m;
    m.Name = str(p[i].name); m.MemoryMB = p[i].memory_mb; m.Core = p[i].core; m.SliceCount = p[i].slice_count;
    for (uint32_t j = 0; j < p[i].n_placements; j++) m.Placements.push_back(MigPlacement{p[i].placements[j].start, p[i].placements[j].size});
    out.push_back(m);
  }
  return out;
}

NodeInfo toNodeInfo(const hs_node_info* n) {
  NodeInfo ni;
  if (!n) return ni;
  ni.ID = str(n->name);
  for (uint32_t i = 0; i < n->n_devices; i++) {
    const hs_device_info& d = n->devices[i];
    DeviceInfo di;
    di.ID = str(d.id); di.Index = d.index; di.Count = d.count; di.Devmem = d.devmem; di.Devcore = d.devcore;
    di.Type = str(d.type); di.Numa = d.numa; di.Mode = str(d.mode); di.Health = d.health != 0;
    di.DeviceVendor = str(d.vendor);
    di.HasNetworkID = d.network_id >= 0; di.NetworkID = d.network_id >= 0 ? d.network_id : 0;
    di.MIGProfiles = toMigProfiles(d.mig_profiles, d.n_mig_profiles);
    if (d.sgpu) { di.HasSgpu = true; di.QosPolicy = str(d.sgpu->qos_policy); di.LinkZone = d.sgpu->link_zone; }
    if (d.enflame) {
      di.EnfMinor = str(d.enflame->minor); di.EnfIndex = str(d.enflame->index); di.EnfMaxSlice = d.enflame->max_slice;
      for (uint32_t k = 0; k < d.enflame->n_profiles; k++) di.EnfProfiles.emplace_back(str(d.enflame->profiles[k].key), str(d.enflame->profiles[k].value));
    }
    if (n->pair_scores)
      for (uint32_t j = 0; j < n->n_devices; j++)
        if (j != i) di.PairScores[str(n->devices[j].id)] = n->pair_scores[i * n->n_devices + j];
    ni.Devices[di.DeviceVendor].push_back(di);
  }
  if (n->device_cordon) ni.DeviceCordon = str(n->device_cordon);
  if (n->hami_vnpu_core >= 0) ni.HamiVnpuCore = n->hami_vnpu_core == 1 ? "true" : "false";
  if (n->metax_losses) ni.MetaxLosses = str(n->metax_losses);
  if (n->metax_scores) ni.MetaxScores = str(n->metax_scores);
  return ni;
}

DeviceListsScore toDls(const oracle_dev& d) {
  DeviceListsScore ds;
  ds.Device.ID = str(d.id); ds.Device.Index = d.index; ds.Device.Used = d.used; ds.Device.Count = d.count;
  ds.Device.Usedmem = d.usedmem; ds.Device.Totalmem = d.totalmem; ds.Device.Totalcore = d.totalcore;
  ds.Device.Usedcores = d.usedcores; ds.Device.Mode = str(d.mode); ds.Device.Numa = d.numa; ds.Device.Type = str(d.type);
  ds.Device.Health = d.health != 0; ds.Device.HasNetworkID = d.network_id >= 0;
  ds.Device.NetworkID = d.network_id >= 0 ? d.network_id : 0;
  ds.Device.MigProfiles = toMigProfiles(d.mig_profiles, d.n_mig_profiles);
  if (d.sgpu) { ds.Device.HasSgpu = true; ds.Device.QosPolicy = str(d.sgpu->qos_policy); ds.Device.LinkZone = d.sgpu->link_zone; }
  ds.Device.OnlinePods = d.online_pods; ds.Device.OfflinePods = d.offline_pods;
  if (d.enflame) {
    ds.Device.EnfMinor = str(d.enflame->minor); ds.Device.EnfIndex = str(d.enflame->index); ds.Device.EnfMaxSlice = d.enflame->max_slice;
    for (uint32_t k = 0; k < d.enflame->n_profiles; k++) ds.Device.EnfProfiles.emplace_back(str(d.enflame->profiles[k].key), str(d.enflame->profiles[k].value));
  }
  for (uint32_t i = 0; d.mig_in_use && i < d.n_mig_in_use; i++)
    ds.Device.MigAllocationsInUse.push_back(MigAllocation{"", MigPlacement{d.mig_in_use[i].start, d.mig_in_use[i].size}});
  ds.Score = d.score;
  return ds;
}

void fillEntries(Arena& a, const PodDevices& pd) {
  for (auto& kv : pd) {
    const char* key = a.s(kv.first);
    for (size_t c = 0; c < kv.second.size(); c++) {
      if (kv.second[c].empty()) {
        a.entries.push_back(hs_alloc_entry{(uint32_t)c, key, nullptr, 0, 0, 0, nullptr, 0, 0});
        continue;
      }
      for (auto& cd : kv.second[c])
        a.entries.push_back(hs_alloc_entry{(uint32_t)c, key, a.s(cd.UUID), cd.Idx, cd.Usedmem, cd.Usedcores,
                                           cd.HasMig ? a.s(cd.MigProfileName) : (cd.HasDrs ? a.s(cd.DrsProfileName) : nullptr),
                                           cd.HasDrs ? (uint32_t)cd.DrsSlice : cd.MigPlace.Start, cd.HasDrs ? 0u : cd.MigPlace.Size});
    }
  }
}

void finish(Arena& a, hs_filter_result* out) {
  out->entries = a.entries.data(); out->n_entries = (uint32_t)a.entries.size();
  out->annotations = a.annos.data(); out->n_annotations = (uint32_t)a.annos.size();
  out->failed = a.failed.data(); out->n_failed = (uint32_t)a.failed.size();
}

}  // namespace

struct oracle_handle {
  std::unique_ptr<Scheduler> s;
  Arena arena;
  std::string err;
};

extern "C" {

oracle_handle* oracle_create(const hs_config* cfg) {
  Config c;
  if (cfg) {
    if (cfg->node_scheduler_policy) c.NodeSchedulerPolicy = cfg->node_scheduler_policy;
    if (cfg->gpu_scheduler_policy) c.GPUSchedulerPolicy = cfg->gpu_scheduler_policy;
    c.AscendHamiVnpuCore = cfg->ascend_hami_vnpu_core != 0;
    for (uint32_t i = 0; i < cfg->n_ascend; i++) {
      AscendConfig ac;
      ac.CommonWord = str(cfg->ascend[i].common_word);
      ac.MemoryAllocatable = cfg->ascend[i].memory_allocatable;
      ac.MemoryCapacity = cfg->ascend[i].memory_capacity;
      for (uint32_t t = 0; t < cfg->ascend[i].n_templates; t++)
        ac.Templates.push_back(AscendTemplate{str(cfg->ascend[i].templates[t].name), cfg->ascend[i].templates[t].memory});
      c.Ascend.push_back(ac);
    }
  }
  if (cfg)
    for (uint32_t i = 0; i < cfg->n_vendors; i++) {
      const hs_vendor_config& v = cfg->vendors[i];
      Config::PlainVendor pv;
      pv.CommonWord = str(v.common_word);
      pv.FitKind = v.fit_kind;
      pv.UseUUID = v.use_uuid_key ? v.use_uuid_key : "hami.io/use-" + pv.CommonWord + "-uuid";            // iluvatar/device.go:71-72
      pv.NoUseUUID = v.nouse_uuid_key ? v.nouse_uuid_key : "hami.io/no-use-" + pv.CommonWord + "-uuid";
      pv.ToAllocate = v.to_allocate_key ? v.to_allocate_key : "hami.io/" + pv.CommonWord + "-devices-to-allocate";
      pv.Allocated = v.allocated_key ? v.allocated_key : "hami.io/" + pv.CommonWord + "-devices-allocated";
      pv.Option = v.option;
      pv.UseType = str(v.use_type_key);
      pv.NoUseType = str(v.nouse_type_key);
      c.Plain.push_back(pv);
    }
  auto* o = new oracle_handle();
  o->s = std::make_unique<Scheduler>(c);
  return o;
}
void oracle_destroy(oracle_handle* o) { delete o; }
void oracle_set_mode(oracle_handle* o, int faithful, int threads) { o->s->faithful = faithful != 0; o->s->threads = threads < 1 ? 1 : threads; }
const char* oracle_last_error(const oracle_handle* o) { return o->err.c_str(); }

int32_t oracle_node_upsert(oracle_handle* o, const hs_node_info* node) {
  try { o->s->AddNode(toNodeInfo(node)); } catch (std::exception& e) { o->err = e.what(); return HS_ERR_INVALID; }
  return HS_OK;
}
int32_t oracle_node_remove(oracle_handle* o, const char* name) { return o->s->RemoveNode(str(name)) ? HS_OK : HS_ERR_NOTFOUND; }
int32_t oracle_usage_apply(oracle_handle* o, const char* pod_uid, const char* node, const char* uuid,
                           int32_t slots, int32_t usedmem, int32_t usedcores) {
  return o->s->UsageApply(str(pod_uid), str(node), str(uuid), slots, usedmem, usedcores) ? HS_OK : HS_ERR_NOTFOUND;
}
// hs_pod_set_mig: the informer re-delivers a cached pod with its decoded reservations (pi.Pod = pod, pods.go:76-77)
int32_t oracle_pod_set_mig(oracle_handle* o, const char* pod_uid, const hs_mig_reservation* res, uint32_t n) {
  const PodInfo* pi = o->s->GetPod(str(pod_uid));
  if (!pi) return HS_ERR_NOTFOUND;
  std::vector<MigReservation> allocs;
  for (uint32_t i = 0; i < n; i++) allocs.push_back(MigReservation{str(res[i].gpu_uuid), str(res[i].profile), MigPlacement{res[i].start, res[i].size}});
  PodInfo copy = *pi;
  return o->s->AddPodWithMig(str(pod_uid), copy.Namespace, copy.NodeID, copy.Devices, allocs) ? HS_OK : HS_ERR_NOTFOUND;
}
int32_t oracle_snapshot_node_mig(oracle_handle* o, const char* node, uint32_t* occupied, uint8_t* health, uint32_t cap) {
  auto snap = o->s->Snapshot(str(node));
  uint32_t n = 0;
  for (auto& d : snap) {
    if (n < cap) {
      uint32_t m = 0;
      for (auto& a : d.MigAllocationsInUse) for (uint32_t b = a.Placement.Start; b < a.Placement.Start + a.Placement.Size && b < 32; b++) m |= 1u << b;
      if (occupied) occupied[n] = m;
      if (health) health[n] = d.Health ? 1 : 0;
    }
    n++;
  }
  return (int32_t)n;
}
/* test hooks: nvidia/mig_topology.go selectMigPlacement / canPlaceMigProfiles and nvidia/device.go selectMigCandidate */
int32_t oracle_mig_select_placement(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                                    const char* profile, hs_mig_placement* out) {
  std::vector<MigPlacement> occ;
  for (uint32_t i = 0; i < n_occ; i++) occ.push_back(MigPlacement{occupied[i].start, occupied[i].size});
  MigPlacement pl;
  if (!SelectMigPlacement(toMigProfiles(profiles, n), occ, str(profile), pl)) return 0;
  if (out) { out->start = pl.Start; out->size = pl.Size; }
  return 1;
}
int32_t oracle_mig_can_place(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                             const char* const* requested, uint32_t n_req) {
  std::vector<MigPlacement> occ;
  for (uint32_t i = 0; i < n_occ; i++) occ.push_back(MigPlacement{occupied[i].start, occupied[i].size});
  std::vector<std::string> req;
  for (uint32_t i = 0; i < n_req; i++) req.push_back(str(requested[i]));
  return CanPlaceMigProfiles(toMigProfiles(profiles, n), occ, req) ? 1 : 0;
}
int32_t oracle_mig_select_candidate(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                                    int32_t memory, int32_t* profile_index, hs_mig_placement* out) {
  std::vector<MigPlacement> occ;
  for (uint32_t i = 0; i < n_occ; i++) occ.push_back(MigPlacement{occupied[i].start, occupied[i].size});
  auto ps = toMigProfiles(profiles, n);
  MigProfile pr; MigPlacement pl;
  if (!SelectMigCandidate(ps, occ, memory, pr, pl)) return 0;
  if (profile_index) { *profile_index = -1; for (uint32_t i = 0; i < n; i++) if (ps[i].Name == pr.Name) { *profile_index = (int32_t)i; break; } }
  if (out) { out->start = pl.Start; out->size = pl.Size; }
  return 1;
}
// MetaxSDevices.ScoreNode (metax/sdevice.go:295-304) on explicit picks: zones[i] = CustomInfo["LinkZone"] of pick i (NULL: no
// CustomInfo), annos = the pod's annotations riding in CustomInfo["Pod.Annotations"], prev = `previous`
float oracle_sgpu_score(oracle_handle* o, const hs_alloc_entry* picks, uint32_t n_picks, const int32_t* zones,
                        const oracle_dev* prev, uint32_t n_prev, const hs_kv* annos, uint32_t n_annos) {
  const Devices* v = o->s->Vendor("Metax-SGPU");
  if (!v) return -1e30f;
  std::map<std::string, std::string> pa;
  for (uint32_t i = 0; i < n_annos; i++) pa[str(annos[i].key)] = str(annos[i].value);
  PodSingleDevice psd;
  for (uint32_t i = 0; i < n_picks; i++) {
    while (psd.size() <= picks[i].container) psd.push_back({});
    ContainerDevice cd;
    cd.UUID = str(picks[i].uuid); cd.Usedcores = picks[i].usedcores; cd.Usedmem = picks[i].usedmem;
    cd.HasSgpu = zones != nullptr; cd.SgpuLinkZone = zones ? zones[i] : 0; cd.PodAnnotations = &pa;
    psd[picks[i].container].push_back(cd);
  }
  std::vector<DeviceUsage> previous;
  for (uint32_t i = 0; i < n_prev; i++) previous.push_back(toDls(prev[i]).Device);
  return v->ScoreNode(psd, "", nullptr, &previous);
}
int32_t oracle_snapshot_node_sgpu(oracle_handle* o, const char* node, int32_t* online, int32_t* offline, const char** pending, uint32_t cap) {
  auto snap = o->s->Snapshot(str(node));
  const Devices* v = o->s->Vendor("Metax-SGPU");
  Arena& a = o->arena;
  a.clear();
  uint32_t n = 0;
  for (auto& d : snap) {
    if (n < cap) {
      if (online) online[n] = d.OnlinePods;
      if (offline) offline[n] = d.OfflinePods;
      std::string q;
      if (pending) pending[n] = (v && d.HasSgpu && v->PendingQos(d.ID, q)) ? a.s(q) : nullptr;
    }
    n++;
  }
  return (int32_t)n;
}
float oracle_kunlun_calc_score(const int32_t* prev, uint32_t n_prev, const int32_t* cur, uint32_t n_cur) {
  auto dedup = [](const int32_t* v, uint32_t n) {       // addidx, kunlun/topo.go:40-48
    std::vector<int> out;
    for (uint32_t i = 0; i < n; i++) if (std::find(out.begin(), out.end(), (int)v[i]) == out.end()) out.push_back((int)v[i]);
    return out;
  };
  return KunlunCalcScore(dedup(prev, n_prev), dedup(cur, n_cur));
}
int32_t oracle_pod_delete(oracle_handle* o, const char* pod_uid) { return o->s->DeletePod(str(pod_uid)) ? HS_OK : HS_ERR_NOTFOUND; }

int32_t oracle_pod_get(oracle_handle* o, const char* pod_uid, const char** node_name, hs_pod_usage_row* rows, uint32_t cap) {
  const PodInfo* pi = o->s->GetPod(str(pod_uid));
  if (!pi) return HS_ERR_NOTFOUND;
  Arena& a = o->arena;
  a.clear();
  if (node_name) *node_name = a.s(pi->NodeID);
  std::vector<std::pair<std::string, std::array<int32_t, 3>>> per;   // by UUID, first-seen order
  for (auto& kv : pi->Devices)
    for (auto& ctr : kv.second)
      for (auto& d : ctr) {
        size_t k = 0;
        while (k < per.size() && per[k].first != d.UUID) k++;
        if (k == per.size()) per.push_back({d.UUID, {0, 0, 0}});
        per[k].second[0] += std::max(d.Slots, 1); per[k].second[1] += d.Usedmem; per[k].second[2] += d.Usedcores;
      }
  uint32_t n = 0;
  for (auto& r : per) {
    if (rows && n < cap) rows[n] = hs_pod_usage_row{a.s(r.first), r.second[0], r.second[1], r.second[2]};
    n++;
  }
  return (int32_t)n;
}

int32_t oracle_filter(oracle_handle* o, const hs_pod* pod, const char* const* node_names, uint32_t n_names,
                      hs_filter_result* out) {
  Arena& a = o->arena;
  a.clear();
  memset(out, 0, sizeof *out);
  out->node_index = -1;
  try {
    Pod p = toPod(pod);
    std::vector<std::string> names;
    if (node_names) for (uint32_t i = 0; i < n_names; i++) names.push_back(str(node_names[i]));
    FilterResult fr = o->s->Filter(p, node_names ? &names : nullptr);
    out->status = fr.status;
    if (fr.status < 0) { o->err = fr.error; return fr.status; }
    if (fr.status == HS_FIT) {
      out->node_index = fr.nodeIndex;
      out->node_name = a.s(fr.chosen.NodeID);
      out->score = fr.chosen.Score;
      fillEntries(a, fr.chosen.Devices);
      for (auto& kv : fr.annotations) a.annos.push_back(hs_kv{a.s(kv.first), a.s(kv.second)});
    }
    // FailedNodes is only returned on no-fit (scheduler.go:1109-1114) but the per-node reasons exist either way
    for (auto& f : fr.failedNodes) a.failed.push_back(hs_failed_node{a.s(f.first), a.s(f.second)});
    finish(a, out);
    return fr.status;
  } catch (std::exception& e) {
    o->err = e.what();
    out->status = HS_ERR_UNSUPPORTED;
    return HS_ERR_UNSUPPORTED;
  }
}

static inline void fnv(uint64_t& hsh, const void* p, size_t n) {
  const unsigned char* b = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) { hsh ^= b[i]; hsh *= 1099511628211ull; }
}
static inline void fnv_str(uint64_t& hsh, const std::string& s) { fnv(hsh, s.c_str(), s.size() + 1); }

int32_t oracle_run_ops(oracle_handle* o, const hs_op* ops, uint32_t n_ops, uint32_t* n_fit, uint32_t* n_nofit,
                       uint64_t* digest, double* seconds) {
  uint64_t hsh = 1469598103934665603ull;
  uint32_t fit = 0, nofit = 0;
  auto t0 = std::chrono::steady_clock::now();
  try {
    for (uint32_t i = 0; i < n_ops; i++) {
      if (ops[i].kind == HS_OP_DELETE) {
        int32_t st = o->s->DeletePod(str(ops[i].pod->uid)) ? HS_OK : HS_ERR_NOTFOUND;
        fnv(hsh, &st, 4);
        continue;
      }
      Pod p = toPod(ops[i].pod);
      FilterResult fr = o->s->Filter(p, nullptr);
      int32_t st = fr.status;
      fnv(hsh, &st, 4);
      if (st == HS_FIT) {
        fit++;
        fnv_str(hsh, fr.chosen.NodeID);
        float sc = fr.chosen.Score;
        fnv(hsh, &sc, 4);
        for (auto& kv : fr.annotations) { fnv_str(hsh, kv.first); fnv_str(hsh, kv.second); }
      } else if (st == HS_NOFIT) {
        nofit++;
        for (auto& f : fr.failedNodes) { fnv_str(hsh, f.first); fnv_str(hsh, f.second); }
      }
    }
  } catch (std::exception& e) {
    o->err = e.what();
    return HS_ERR_UNSUPPORTED;
  }
  auto t1 = std::chrono::steady_clock::now();
  if (n_fit) *n_fit = fit;
  if (n_nofit) *n_nofit = nofit;
  if (digest) *digest = hsh;
  if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
  return HS_OK;
}

int32_t oracle_snapshot_node(oracle_handle* o, const char* node, hs_device_usage_row* rows, uint32_t cap) {
  auto v = o->s->Snapshot(str(node));
  uint32_t n = 0;
  for (auto& d : v) { if (n >= cap) break; rows[n++] = hs_device_usage_row{d.Used, d.Usedmem, d.Usedcores}; }
  return (int32_t)v.size();
}

static NodeUsage makeUsage(oracle_handle* o, const oracle_dev* devs, uint32_t n, const char* list_policy,
                           int32_t numa_bind, const Pod& pod, const NodeInfo* ni) {
  NodeUsage nu;
  nu.Info = ni;
  nu.Devices.Policy = list_policy ? std::string(list_policy) : GetGPUSchedulerPolicyByPod(o->s->config().GPUSchedulerPolicy, pod);
  if (numa_bind >= 0) nu.Devices.NumaBind = numa_bind != 0;
  else {
    auto it = pod.Annotations.find("nvidia.com/numa-bind");
    nu.Devices.NumaBind = it != pod.Annotations.end() &&
                          (it->second == "1" || it->second == "t" || it->second == "T" || it->second == "TRUE" ||
                           it->second == "true" || it->second == "True");
  }
  for (uint32_t i = 0; i < n; i++) nu.Devices.DeviceLists.push_back(toDls(devs[i]));
  return nu;
}

int32_t oracle_score_node(oracle_handle* o, const char* node_id, const oracle_dev* devs, uint32_t n,
                          const char* list_policy, int32_t numa_bind, const hs_pod* pod,
                          const hs_node_info* node_info, hs_filter_result* out) {
  Arena& a = o->arena;
  a.clear();
  memset(out, 0, sizeof *out);
  try {
    Pod p = toPod(pod);
    NodeInfo ni = toNodeInfo(node_info);
    if (!node_info) ni.ID = str(node_id);
    NodeUsage nu = makeUsage(o, devs, n, list_policy, numa_bind, p, &ni);
    DeviceScoringWeights w;
    auto wit = p.Annotations.find("hami.io/device-scoring-weights");
    if (wit != p.Annotations.end()) {
      std::string err;
      if (!ParseDeviceScoringWeights(wit->second, w, err)) { o->err = err; out->status = HS_ERR_WEIGHTS; return HS_ERR_WEIGHTS; }
    }
    std::string nodePolicy = o->s->config().NodeSchedulerPolicy;
    auto nit = p.Annotations.find("hami.io/node-scheduler-policy");
    if (nit != p.Annotations.end()) nodePolicy = nit->second;
    NodeScoreResult r = o->s->ScoreNode(str(node_id), nu, p, nodePolicy, w);
    if (r.fit) {
      out->status = HS_FIT;
      out->node_name = a.s(str(node_id));
      out->score = r.score.Score;
      fillEntries(a, r.score.Devices);
    } else {
      out->status = HS_NOFIT;
      a.failed.push_back(hs_failed_node{a.s(str(node_id)), a.s(r.reason)});
    }
    finish(a, out);
    return out->status;
  } catch (std::exception& e) {
    o->err = e.what();
    out->status = HS_ERR_UNSUPPORTED;
    return HS_ERR_UNSUPPORTED;
  }
}

int32_t oracle_fit_in_devices(oracle_handle* o, const oracle_dev* devs, uint32_t n, const char* list_policy,
                              int32_t numa_bind, const hs_container* ctr, const hs_pod* pod,
                              const hs_node_info* node_info, hs_filter_result* out) {
  Arena& a = o->arena;
  a.clear();
  memset(out, 0, sizeof *out);
  try {
    Pod p = toPod(pod);
    NodeInfo ni = toNodeInfo(node_info);
    NodeUsage nu = makeUsage(o, devs, n, list_policy, numa_bind, p, &ni);
    ContainerDeviceRequests reqs;
    for (uint32_t r = 0; r < ctr->n_reqs; r++) {
      const hs_container_request& q = ctr->reqs[r];
      ContainerDeviceRequest k;
      k.Nums = q.nums; k.Type = str(q.type); k.Memreq = q.memreq; k.MemPercentagereq = q.mempct; k.Coresreq = q.coresreq;
      reqs[str(q.key)] = k;
    }
    DeviceScoringWeights w;
    auto wit = p.Annotations.find("hami.io/device-scoring-weights");
    if (wit != p.Annotations.end()) { std::string err; ParseDeviceScoringWeights(wit->second, w, err); }
    PodDevices pd;
    std::string reason;
    bool fit = o->s->FitInDevices(nu, reqs, p, pd, w, reason);
    out->status = fit ? HS_FIT : HS_NOFIT;
    fillEntries(a, pd);
    if (!fit) a.failed.push_back(hs_failed_node{a.s(ni.ID), a.s(reason)});
    finish(a, out);
    return fit ? 1 : 0;
  } catch (std::exception& e) {
    o->err = e.what();
    out->status = HS_ERR_UNSUPPORTED;
    return HS_ERR_UNSUPPORTED;
  }
}

int32_t oracle_fit(oracle_handle* o, const char* vendor_key, const oracle_dev* devs, uint32_t n,
                   const hs_container_request* req, const hs_pod* pod, const hs_node_info* node_info,
                   hs_filter_result* out) {
  Arena& a = o->arena;
  a.clear();
  memset(out, 0, sizeof *out);
  try {
    const Devices* v = o->s->Vendor(str(vendor_key));
    if (!v) { o->err = "unknown vendor"; return HS_ERR_NOTFOUND; }
    Pod p = toPod(pod);
    NodeInfo ni = toNodeInfo(node_info);
    std::vector<DeviceUsage> store;
    for (uint32_t i = 0; i < n; i++) store.push_back(toDls(devs[i]).Device);
    std::vector<DeviceUsage*> ptrs;
    for (auto& d : store) ptrs.push_back(&d);
    ContainerDeviceRequest k;
    k.Nums = req->nums; k.Type = str(req->type); k.Memreq = req->memreq; k.MemPercentagereq = req->mempct; k.Coresreq = req->coresreq;
    FitResult fr = v->Fit(ptrs, k, p, node_info ? &ni : nullptr, o->s->config().GPUSchedulerPolicy);
    out->status = fr.fit ? HS_FIT : HS_NOFIT;
    const char* key = a.s(k.Type);
    for (auto& cd : fr.devs) a.entries.push_back(hs_alloc_entry{0, key, a.s(cd.UUID), cd.Idx, cd.Usedmem, cd.Usedcores,
                                                               cd.HasDrs ? a.s(cd.DrsProfileName) : nullptr, cd.HasDrs ? (uint32_t)cd.DrsSlice : 0u, 0});
    a.failed.push_back(hs_failed_node{a.s(""), a.s(fr.reason)});
    finish(a, out);
    return fr.fit ? 1 : 0;
  } catch (std::exception& e) {
    o->err = e.what();
    out->status = HS_ERR_UNSUPPORTED;
    return HS_ERR_UNSUPPORTED;
  }
}

float oracle_compute_score(const oracle_dev* d, const hs_container_request* reqs, uint32_t n,
                           int64_t w_slot, int64_t w_core, int64_t w_mem) {
  DeviceListsScore ds = toDls(*d);
  ContainerDeviceRequests m;
  for (uint32_t r = 0; r < n; r++) {
    ContainerDeviceRequest k;
    k.Nums = reqs[r].nums; k.Type = str(reqs[r].type); k.Memreq = reqs[r].memreq; k.MemPercentagereq = reqs[r].mempct; k.Coresreq = reqs[r].coresreq;
    m[str(reqs[r].key)] = k;
  }
  ds.ComputeScore(m, DeviceScoringWeights{w_slot, w_core, w_mem});
  return ds.Score;
}

int32_t oracle_less(const char* policy, int32_t numa_bind, const oracle_dev* a, const oracle_dev* b) {
  DeviceUsageList l;
  l.Policy = str(policy);
  l.NumaBind = numa_bind != 0;
  l.DeviceLists.push_back(toDls(*a));
  l.DeviceLists.push_back(toDls(*b));
  return l.Less(0, 1) ? 1 : 0;
}

void oracle_sort(const char* policy, int32_t numa_bind, const oracle_dev* devs, uint32_t n, int32_t* order_out) {
  DeviceUsageList l;
  l.Policy = str(policy);
  l.NumaBind = numa_bind != 0;
  for (uint32_t i = 0; i < n; i++) {
    DeviceListsScore ds = toDls(devs[i]);
    ds.Device.Usedmem = ds.Device.Usedmem;  // keep
    l.DeviceLists.push_back(ds);
  }
  // remember original positions through a parallel tag: reuse NetworkID (not read by Less)
  for (uint32_t i = 0; i < n; i++) l.DeviceLists[i].Device.NetworkID = (int)i;
  l.Sort();
  for (uint32_t i = 0; i < n; i++) order_out[i] = l.DeviceLists[i].Device.NetworkID;
}

float oracle_node_default_score(const oracle_dev* devs, uint32_t n) {
  DeviceUsageList l;
  for (uint32_t i = 0; i < n; i++) l.DeviceLists.push_back(toDls(devs[i]));
  return ComputeDefaultScore(l);
}

static std::map<std::string, std::string> toMap(const hs_kv* annos, uint32_t n) {
  std::map<std::string, std::string> m;
  for (uint32_t i = 0; i < n; i++) m[str(annos[i].key)] = str(annos[i].value);
  return m;
}
int32_t oracle_check_uuid(const hs_kv* annos, uint32_t n, const char* id, const char* use_key, const char* nouse_key) {
  return CheckUUID(toMap(annos, n), str(id), str(use_key), str(nouse_key)) ? 1 : 0;
}
int32_t oracle_check_type(const hs_kv* annos, uint32_t n, const char* card_type, const char* use_key, const char* nouse_key) {
  return CheckType(toMap(annos, n), str(card_type), str(use_key), str(nouse_key)) ? 1 : 0;
}
int32_t oracle_parse_weights(const char* value, int64_t out3[3]) {
  DeviceScoringWeights w;
  std::string err;
  if (!ParseDeviceScoringWeights(str(value), w, err)) return HS_ERR_WEIGHTS;
  out3[0] = w.Slot; out3[1] = w.Core; out3[2] = w.Memory;
  return HS_OK;
}
int32_t oracle_gen_reason(const int32_t* counts, int32_t cards, char* buf, size_t cap) {
  std::map<std::string, int> m;
  for (int i = 0; i < HS_R_COUNT; i++) if (counts[i]) m[kReasonNames[i]] = counts[i];
  std::string s = GenReason(m, cards);
  snprintf(buf, cap, "%s", s.c_str());
  return (int32_t)s.size();
}
int32_t oracle_collapse(const hs_alloc_entry* in, uint32_t n_in, uint32_t n_init, hs_alloc_entry* out, uint32_t cap) {
  Arena& a = g_arena;
  a.clear();
  PodDevices raw;
  for (uint32_t i = 0; i < n_in; i++) {
    PodSingleDevice& psd = raw[str(in[i].type_key)];
    if (psd.size() <= in[i].container) psd.resize(in[i].container + 1);
    if (!in[i].uuid) continue;
    ContainerDevice cd;
    cd.UUID = str(in[i].uuid); cd.Type = str(in[i].type_key); cd.Usedmem = in[i].usedmem; cd.Usedcores = in[i].usedcores;
    psd[in[i].container].push_back(cd);
  }
  PodDevices c = CollapseInitContainerUsage((int)n_init, raw);
  uint32_t n = 0;
  for (auto& kv : c)
    for (auto& cds : kv.second)
      for (auto& cd : cds) {
        if (n < cap) out[n] = hs_alloc_entry{0, a.s(kv.first), a.s(cd.UUID), cd.Slots, cd.Usedmem, cd.Usedcores, nullptr, 0, 0};
        n++;
      }
  return (int32_t)n;
}

static std::string quota_name(oracle_handle* o, const char* vendor, int32_t which) {
  const Devices* v = o->s->Vendor(str(vendor));
  if (!v) return "";
  ResourceNames rn = v->GetResourceNames();
  return which == 0 ? rn.Memory : rn.Core;
}
int32_t oracle_quota_set(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t limit) {
  std::string name = quota_name(o, vendor, which);
  if (name.empty()) return HS_ERR_INVALID;    /* IsManagedQuota false: not tracked */
  o->s->quota.SetLimit(str(ns), name, limit);
  return HS_OK;
}
int32_t oracle_quota_del(oracle_handle* o, const char* ns, const char* vendor, int32_t which) {
  std::string name = quota_name(o, vendor, which);
  if (name.empty()) return HS_ERR_INVALID;
  o->s->quota.ClearLimit(str(ns), name);
  return HS_OK;
}
int32_t oracle_quota_get(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t* limit, int64_t* used, int32_t* limit_set) {
  std::string name = quota_name(o, vendor, which);
  auto dq = o->s->quota.Quotas.find(str(ns));
  if (name.empty() || dq == o->s->quota.Quotas.end()) return HS_ERR_NOTFOUND;
  auto q = dq->second.find(name);
  if (q == dq->second.end()) return HS_ERR_NOTFOUND;
  if (limit) *limit = q->second.Limit;
  if (used) *used = q->second.Used;
  if (limit_set) *limit_set = q->second.LimitSet ? 1 : 0;
  return HS_OK;
}
int32_t oracle_quota_add_usage(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t delta) {
  std::string name = quota_name(o, vendor, which);
  if (name.empty()) return HS_ERR_INVALID;
  std::map<std::string, int64_t> u{{name, delta < 0 ? -delta : delta}};
  if (delta < 0) o->s->quota.RmUsage(str(ns), u); else o->s->quota.AddUsage(str(ns), u);
  return HS_OK;
}
float oracle_vendor_score_node(oracle_handle* o, const char* vendor, uint32_t n_picks, const char* policy, const char* metax_losses,
                               const char* metax_scores) {
  const Devices* v = o->s->Vendor(str(vendor));
  if (!v) return 0;
  NodeInfo ni;
  if (metax_losses) ni.MetaxLosses = str(metax_losses);
  if (metax_scores) ni.MetaxScores = str(metax_scores);
  PodSingleDevice psd(1);
  for (uint32_t i = 0; i < n_picks; i++) { ContainerDevice cd; cd.Idx = (int)i; cd.UUID = "d" + std::to_string(i); cd.Type = str(vendor); psd[0].push_back(cd); }
  return v->ScoreNode(psd, str(policy), &ni);
}
int32_t oracle_fit_quota_scalar(oracle_handle* o, const char* ns, const char* vendor, int64_t memreq, int32_t factor, int64_t coresreq) {
  const Devices* v = o->s->Vendor(str(vendor));
  if (!v) return 1;                       /* FitQuota: device not present → true (quota.go:63-66) */
  ResourceNames rn = v->GetResourceNames();
  rn.MemoryFactor = factor;
  return o->s->quota.FitQuota(str(ns), memreq, rn, coresreq) ? 1 : 0;
}
int32_t oracle_fit_quota(oracle_handle* o, const char* ns, uint32_t n_init, const hs_alloc_entry* tmp, uint32_t n_tmp,
                         const hs_alloc_entry* allocated, uint32_t n_allocated, int32_t has_allocated, const char* vendor,
                         const char* dev_uuid, int64_t memreq, int64_t coresreq) {
  Pod pod;
  pod.Namespace = str(ns);
  pod.NumInitContainers = (int)n_init;
  ContainerDevices t;
  for (uint32_t i = 0; i < n_tmp; i++) {
    ContainerDevice cd;
    cd.UUID = str(tmp[i].uuid); cd.Type = str(tmp[i].type_key); cd.Usedmem = tmp[i].usedmem; cd.Usedcores = tmp[i].usedcores;
    t.push_back(cd);
  }
  PodDevices alloc;
  for (uint32_t i = 0; i < n_allocated; i++) {
    PodSingleDevice& psd = alloc[str(allocated[i].type_key)];
    if (psd.size() <= allocated[i].container) psd.resize(allocated[i].container + 1);
    if (!allocated[i].uuid) continue;
    ContainerDevice cd;
    cd.UUID = str(allocated[i].uuid); cd.Type = str(allocated[i].type_key); cd.Usedmem = allocated[i].usedmem; cd.Usedcores = allocated[i].usedcores;
    psd[allocated[i].container].push_back(cd);
  }
  return FitQuotaForDevice(*o->s, pod, t, has_allocated ? &alloc : nullptr, str(vendor), str(dev_uuid), memreq, coresreq) ? 1 : 0;
}

}  // extern "C"
