// hs_eval.cuh — per-node evaluation of one pod on the device: the integer-level
// mirror of scoreNode → fitInDevices → <vendor>.Fit → ComputeScore/Less →
// ComputeDefaultScore/OverrideScore (reference: pkg/scheduler/score.go:81-346,
// pkg/scheduler/policy/*.go, pkg/device/{nvidia,hygon,cambricon,ascend}/device.go Fit).
//
// One thread evaluates one node.  All float math is IEEE binary32 with every
// operation rounded separately, in the reference's evaluation order (the file
// is compiled with -fmad=false and uses __fdiv_rn/__fmul_rn/__fadd_rn), so
// scores and therefore tie-breaks are bit-identical to Go on amd64.
#pragma once
#include "hs_types.cuh"

// Working set of one node while a pod is fitted on it.  Arrays are indexed with
// run-time device slots (sorted order), so they live in local memory / L1.
struct NodeWork {
  int used[HS_D], mem[HS_D], cores[HS_D];            // mutable usage copy (NodeUsage.DeepCopy, score.go:320)
  int count[HS_D], totmem[HS_D], totcore[HS_D];      // capacity
  uint32_t packed[HS_D], dflags[HS_D];
  float score[HS_D];                                 // DeviceListsScore.Score
  uint32_t order;                                    // nibble i = device slot at list position i
  uint32_t mth_prev;                                 // EF_MTHREADS: device slots the pod's earlier containers (of this phase) received
  uint32_t sg_picks, sg_excl;                        // EF_SGPU: device slots the pod received so far (mask) / how often each was received whole
                                                     // (Usedcores 100; 4 bits per slot: ScoreNode's allocatedDevices is a LIST), all containers
  uint32_t kl_cur;                                   // EF_KUNLUN: card indices the pod received so far, all containers (ScoreNode's `current`)
  int aws[HS_D];                                     // EF_NEURON: CustomInfo["awsusageinfo"] of the working copy, -1 = key absent
  int ndev;
};

enum { EV_SCORE = 0, EV_RECORD = 1, EV_EXPLAIN = 2 };

struct EvalOut {
  // EV_RECORD
  DevAlloc* allocs;
  uint32_t n_alloc;
  // EV_EXPLAIN
  uint32_t fkind, fcards;
  uint32_t reasons[24];
};

__device__ __forceinline__ int ord_get(uint32_t order, int i) { return (order >> (4 * i)) & 7; }
__device__ __forceinline__ uint32_t ord_swap(uint32_t order, int i, int j) {
  uint32_t a = (order >> (4 * i)) & 15u, b = (order >> (4 * j)) & 15u;
  order &= ~((15u << (4 * i)) | (15u << (4 * j)));
  return order | (b << (4 * i)) | (a << (4 * j));
}

// DeviceListsScore.ComputeScore, policy/gpu_policy.go:194-222
__device__ __forceinline__ float dev_compute_score(const NodeWork& w, int d, const DevPod& pod, const DevEntry* ent, int eb, int ee) {
  if (w.count[d] == 0 || w.totcore[d] == 0 || w.totmem[d] == 0) return 0.0f;
  int request = 0, core = 0, mem = 0;
  const uint64_t tbit = 1ull << PK_TYPE(w.packed[d]);
  for (int e = eb; e < ee; e++) {
    if (!(ent[e].type_equal & tbit)) continue;
    request += 1;
    core += ent[e].coresreq;
    if (ent[e].mempct != 0 && ent[e].mempct != 101) {
      mem += (int)(((long long)w.totmem[d] * (long long)ent[e].mempct) / 100);
      continue;
    }
    mem += ent[e].memreq;
  }
  float usedScore = __fdiv_rn((float)(request + w.used[d]), (float)w.count[d]);
  float coreScore = __fdiv_rn((float)(core + w.cores[d]), (float)w.totcore[d]);
  float memScore = __fdiv_rn((float)(mem + w.mem[d]), (float)w.totmem[d]);
  float s = __fadd_rn(__fmul_rn(pod.w_slot, usedScore), __fmul_rn(pod.w_core, coreScore));
  s = __fadd_rn(s, __fmul_rn(pod.w_mem, memScore));
  return __fmul_rn(10.0f, s);
}

// DeviceUsageList.Less, policy/gpu_policy.go:77-167 (a, b are device slots)
__device__ __forceinline__ bool dev_less(const NodeWork& w, const DevPod& pod, int a, int b) {
  const float si = w.score[a], sj = w.score[b];
  const int ni = PK_NUMA(w.packed[a]), nj = PK_NUMA(w.packed[b]);
  switch (pod.sort_mode) {
    case SORT_CHAIN: {
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int key = pod.chain[k];
        if (key == CK_BINPACK) { if (si != sj) return si < sj; }
        else if (key == CK_SPREAD) { if (si != sj) return si > sj; }
        else if (key == CK_NUMA) { if (ni != nj) return ni < nj; }
      }
      return PK_INDEX(w.packed[a]) < PK_INDEX(w.packed[b]);
    }
    case SORT_MUTEX: {
      const int ui = w.used[a], uj = w.used[b];
      if (ui != uj) return ui > uj;
      return ni < nj;
    }
    case SORT_NUMA_BINPACK:
      if (ni == nj) return si < sj;
      return ni > nj;
    case SORT_NUMA_SPREAD:
      if (ni == nj) return si > sj;
      return ni < nj;
    case SORT_BINPACK:
      if (si != sj) return si < sj;
      return ni < nj;
    default:  // SORT_SPREAD
      if (si != sj) return si > sj;
      return ni < nj;
  }
}

// sort.Sort for n <= 12 == stable insertion sort (Go stdlib pdqsort, maxInsertion 12)
__device__ __forceinline__ void dev_sort(NodeWork& w, const DevPod& pod) {
  uint32_t order = w.order;
  for (int i = 1; i < w.ndev; i++) {
    for (int j = i; j > 0 && dev_less(w, pod, ord_get(order, j), ord_get(order, j - 1)); j--)
      order = ord_swap(order, j, j - 1);
  }
  w.order = order;
}

// NodeScore.ComputeDefaultScore, policy/node_policy.go:106-130
__device__ __forceinline__ float node_default_score(const NodeWork& w) {
  int used = 0, usedCore = 0, usedMem = 0, total = 0, totalCore = 0, totalMem = 0;
  for (int d = 0; d < w.ndev; d++) {
    used += w.used[d]; usedCore += w.cores[d]; usedMem += w.mem[d];
    total += w.count[d]; totalCore += w.totcore[d]; totalMem += w.totmem[d];
  }
  if (total == 0 || totalCore == 0 || totalMem == 0) return 0.0f;
  float useScore = __fdiv_rn((float)used, (float)total);
  float coreScore = __fdiv_rn((float)usedCore, (float)totalCore);
  float memScore = __fdiv_rn((float)usedMem, (float)totalMem);
  return __fmul_rn(10.0f, __fadd_rn(__fadd_rn(useScore, coreScore), memScore));
}

// use/no-use UUID lists resolved by the host into sorted (node<<8 | mask) words
__device__ __forceinline__ uint32_t uuid_mask_lookup(const uint32_t* pool, uint32_t off, uint32_t cnt, uint32_t node) {
  uint32_t lo = 0, hi = cnt;
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    uint32_t v = pool[off + mid];
    if ((v >> 8) < node) lo = mid + 1; else hi = mid;
  }
  if (lo < cnt) {
    uint32_t v = pool[off + lo];
    if ((v >> 8) == node) return v & 255u;
  }
  return 0;
}

struct FitCtx {
  int tmp_dev[HS_D];    // tmpDevs[k.Type]: device slots in pick order
  int tmp_mem[HS_D];    // their resolved memreq
  int tmp_cores[HS_D];  // EF_AMD / EF_NEURON: the per-device Usedcores (every other vendor records `cores` for all picks)
  int tmp_aws[HS_D];    // EF_NEURON: the pick's CustomInfo["awsusageinfo"]
  int ntmp;
  int cores;            // k.Coresreq after the 100 clamp
};

// Namespace quota (device/quota.go) as the NVIDIA and MLU Fit see it.  fitQuota (nvidia/device.go:666-692,
// cambricon/device.go:351-376) collapses `allocated` — the containers of this pod placed so far on this node, as the
// caller built the list — plus the current container (picks so far + the candidate) with CollapseInitContainerUsage
// and checks the sum.  A list position below numInit counts as an init container (peak per device), the rest as app
// containers (sum per device); positions restart at 0 for the app phase because score.Devices starts empty
// (score.go:323-333), which this restates as is.  Index 0 = key "NVIDIA", 1 = key "MLU".
struct QuotaAcc {
  long long lim_mem[2], used_mem[2], lim_core[2], used_core[2];
  bool on[2];
  int n_init;
  int len[2];                                // containers in allocated[key], base-type padding included
  int ipm[2][HS_D], ipc[2][HS_D];            // per device: init peak memory / cores
  int apm[2][HS_D], apc[2][HS_D];            // per device: app sum memory / cores
};
__device__ __forceinline__ void quota_reset_lists(QuotaAcc& q) {
  for (int v = 0; v < 2; v++) {
    q.len[v] = 0;
    for (int d = 0; d < HS_D; d++) { q.ipm[v][d] = q.ipc[v][d] = q.apm[v][d] = q.apc[v][d] = 0; }
  }
}
__device__ __forceinline__ void quota_load(QuotaAcc& q, const DevPod& pod, const KernelArgs& ka) {
  q.n_init = pod.n_init;
  for (int v = 0; v < 2; v++) {
    q.on[v] = pod.q_idx[v] >= 0;
    if (q.on[v]) {
      const long long* r = ka.quota + (size_t)pod.q_idx[v] * 4;
      // through L2: another CTA's commit changed `used` since this SM last read the row
      q.lim_mem[v] = __ldcg(r + 0); q.used_mem[v] = __ldcg(r + 1); q.lim_core[v] = __ldcg(r + 2); q.used_core[v] = __ldcg(r + 3);
    }
  }
  quota_reset_lists(q);
}
// would the pod's collapsed total stay inside the quota if device d joined the current container's picks?
__device__ __forceinline__ bool quota_fits(const QuotaAcc& q, int v, const int* tmp_dev, const int* tmp_mem, int ntmp,
                                           int d, int memreq, int cores) {
  const bool as_init = q.len[v] < q.n_init;
  long long mem = 0, core = 0;
  for (int dd = 0; dd < HS_D; dd++) {
    int cm = 0, cc = 0;
    bool cur = dd == d;
    if (cur) { cm = memreq; cc = cores; }
    for (int t = 0; t < ntmp; t++)
      if (tmp_dev[t] == dd) { cm += tmp_mem[t]; cc += cores; cur = true; }
    int im = q.ipm[v][dd], ic = q.ipc[v][dd], am = q.apm[v][dd], ac = q.apc[v][dd];
    if (cur) {
      if (as_init) { im = cm > im ? cm : im; ic = cc > ic ? cc : ic; }
      else { am += cm; ac += cc; }
    }
    mem += im > am ? im : am;
    core += ic > ac ? ic : ac;
  }
  if (q.used_mem[v] + mem > q.lim_mem[v]) return false;
  if (q.used_core[v] + core > q.lim_core[v]) return false;
  return true;
}
// devinput[k.Type] = append(devinput[k.Type], picks) (score.go:132)
__device__ __forceinline__ void quota_append(QuotaAcc& q, int v, const int* tmp_dev, const int* tmp_mem, int ntmp, int cores) {
  const bool as_init = q.len[v] < q.n_init;
  for (int t = 0; t < ntmp; t++) {
    const int d = tmp_dev[t];
    if (as_init) {
      if (tmp_mem[t] > q.ipm[v][d]) q.ipm[v][d] = tmp_mem[t];
      if (cores > q.ipc[v][d]) q.ipc[v][d] = cores;
    } else {
      q.apm[v][d] += tmp_mem[t];
      q.apc[v][d] += cores;
    }
  }
  q.len[v]++;
}

#define REASON(r) do { if (MODE == EV_EXPLAIN) out.reasons[r]++; } while (0)

// <vendor>.Fit: nvidia/device.go:715-876, hygon/device.go:264-369, cambricon/device.go:378-488,
// ascend/device.go:442-617 — one backward scan with per-vendor gates.
// td[0..cnt) = typeDevices (device slots, sorted list order).  Returns fit; the picks are in fc.
template <int MODE>
__device__ bool dev_fit(NodeWork& w, const DevPod& pod, const DevEntry& e, const int* td, int cnt,
                        uint32_t node, uint32_t nflags, const KernelArgs& ka, FitCtx& fc, EvalOut& out, const QuotaAcc* qa) {
  const int vendor = e.vendor;
  const int qv = vendor == HV_NVIDIA ? 0 : (vendor == HV_MLU ? 1 : -1);
  const bool quota_on = qa != nullptr && qv >= 0 && qa->on[qv];
  const bool isNv = vendor == HV_NVIDIA;
  const bool isAscend = vendor >= HV_ASCEND0;
  int k_nums = e.nums;
  const int origin = e.nums;
  int prevnuma = -1;
  fc.ntmp = 0;
  const bool countOnly = e.flags & EF_COUNT_ONLY;
  fc.cores = (e.coresreq > 100 && !countOnly && !(e.flags & EF_WHOLE_CARD)) ? 100 : e.coresreq;   // biren / GCU record Coresreq as asked
  const bool isMutex = pod.flags & PF_MUTEX;
  bool needTopology = isNv && (pod.flags & PF_TOPOLOGY);
  const bool isHAMiCore = isAscend && (pod.flags & PF_ASCEND_HAMICORE);
  if (MODE == EV_EXPLAIN) {
    for (int r = 0; r < 24; r++) out.reasons[r] = 0;
    out.fkind = FK_HIST;
    out.fcards = (uint32_t)cnt;
  }
  if (e.flags & EF_SGPU) {
    // metax sGPU, metax/sdevice.go:332-495: filter every device, then prioritise among those that passed
    const int cls = pod.sg_class & 3;
    const bool online = cls == 1, reqClassed = cls == 1 || cls == 2;
    const uint32_t reqq = pod.sg_qos;
    uint32_t useMaskS = 0xffu, nouseMaskS = 0u;
    if (e.has_use) useMaskS = uuid_mask_lookup(ka.uuid_pool, e.use_off, e.use_cnt, node);
    if (e.nouse_cnt) nouseMaskS = uuid_mask_lookup(ka.uuid_pool, e.nouse_off, e.nouse_cnt, node);
    int cand[HS_D], nc = 0;
    for (int i = cnt - 1; i >= 0; i--) {
      const int d = td[i];
      const uint32_t df = w.dflags[d];
      if (!(df & DF_HEALTH)) { REASON(15 /*CardUnhealthy*/); continue; }
      if (w.used[d] != 0) {                                           // checkAppClass, :664-703
        const bool devOn = SG_ON(df) > 0, devOff = SG_OFF(df) > 0;
        if (((devOff || devOn) != reqClassed) || (devOn && cls == 2)) { REASON(16 /*CardAppClassMismatch*/); continue; }
      }
      if (!online && w.used[d] != 0 && e.coresreq != 100) {           // checkDeviceQos, :527-556: a switch in flight wins over the device's own policy
        const uint32_t devq = SG_JQ(df) ? SG_JQID(df) : DF_MODE(df);
        if (devq != 0u && devq != reqq) { REASON(17 /*CardQosPolicyMismatch*/); continue; }
      }
      if (!((useMaskS >> d) & 1u) || ((nouseMaskS >> d) & 1u)) { REASON(1); continue; }
      if (w.count[d] <= w.used[d]) { REASON(2); continue; }
      if (isMutex && w.used[d] > 0) { REASON(9); continue; }
      const int memreq = e.memreq > 0 ? e.memreq : (int)((unsigned)w.totmem[d] * (unsigned)e.mempct) / 100;
      if (w.totmem[d] - w.mem[d] < memreq) { REASON(4); continue; }
      if (!online) {
        if (w.totcore[d] - w.cores[d] < e.coresreq) { REASON(5); continue; }
        if (w.totcore[d] == 100 && e.coresreq == 100 && w.used[d] > 0) { REASON(9); continue; }
        if (w.totcore[d] != 0 && w.cores[d] == w.totcore[d] && e.coresreq == 0) { REASON(3); continue; }
      }
      cand[nc++] = d;
    }
    const int nums = e.nums < 0 ? 0 : e.nums;
    if (nc < nums) {
      if (MODE == EV_EXPLAIN && nc > 0) out.reasons[12] = (uint32_t)nc;
      return false;
    }
    if (nc > nums) {
      if (cls != 0) {
        if (online) {
          // prioritizeOnlineDevices, :717-770: Go's insertion sort (<= 12 elements) with the online / offline comparator
          for (int i = 1; i < nc; i++)
            for (int j = i; j > 0; j--) {
              const uint32_t a = w.dflags[cand[j]], b = w.dflags[cand[j - 1]];
              const int ao = (int)SG_ON(a), af = (int)SG_OFF(a), bo = (int)SG_ON(b), bf_ = (int)SG_OFF(b);
              const bool aOnly = ao == 0 && af > 0, bOnly = bo == 0 && bf_ > 0;
              bool less;
              if (aOnly && bOnly) less = af < bf_;
              else if (aOnly) less = false;
              else if (bOnly) less = true;
              else less = ao == bo ? af < bf_ : ao < bo;
              if (!less) break;
              const int t = cand[j]; cand[j] = cand[j - 1]; cand[j - 1] = t;
            }
        }
      } else if (e.coresreq == 100) {
        // prioritizeExclusiveDevices, :583-645: MetaLink zones in ascending id (canonical rule for the Go map), ordered by
        // size with the same insertion sort; a zone that holds enough cards, else across zones, else unlinked cards
        int zone_of[HS_D], zsize[HS_D], zones[HS_D], nz = 0;
        for (int i = 0; i < nc; i++) {
          const uint32_t df = w.dflags[cand[i]];
          zone_of[i] = (df & DF_SGPU) ? (int)DF_NET(df) : 0;
          if (zone_of[i] == 255) zone_of[i] = 0;
        }
        for (int i = 0; i < nc; i++) {
          if (zone_of[i] <= 0) continue;
          bool seen = false;
          for (int z = 0; z < nz; z++) if (zones[z] == zone_of[i]) { zsize[z]++; seen = true; }
          if (!seen) { zones[nz] = zone_of[i]; zsize[nz] = 1; nz++; }
        }
        for (int i = 1; i < nz; i++)                 // ascending zone id
          for (int j = i; j > 0 && zones[j] < zones[j - 1]; j--) { int t = zones[j]; zones[j] = zones[j - 1]; zones[j - 1] = t; t = zsize[j]; zsize[j] = zsize[j - 1]; zsize[j - 1] = t; }
        for (int i = 1; i < nz; i++)                 // then by size, stable
          for (int j = i; j > 0 && zsize[j] < zsize[j - 1]; j--) { int t = zones[j]; zones[j] = zones[j - 1]; zones[j - 1] = t; t = zsize[j]; zsize[j] = zsize[j - 1]; zsize[j - 1] = t; }
        int best[HS_D], nb = 0;
        for (int z = 0; z < nz && nb == 0; z++)
          if (zsize[z] >= nums)
            for (int i = 0; i < nc && nb < nums; i++) if (zone_of[i] == zones[z]) best[nb++] = cand[i];
        if (nb < nums) {
          nb = 0;
          for (int z = 0; z < nz && nb < nums; z++)
            for (int i = 0; i < nc && nb < nums; i++) if (zone_of[i] == zones[z]) best[nb++] = cand[i];
          for (int i = 0; i < nc && nb < nums; i++) if (zone_of[i] <= 0) best[nb++] = cand[i];
        }
        if (nb >= nums) for (int i = 0; i < nums; i++) cand[i] = best[i];
      }
    }
    for (int t = 0; t < nums && t < HS_D; t++) {
      const int d = cand[t];
      fc.tmp_dev[fc.ntmp] = d;
      fc.tmp_mem[fc.ntmp] = e.memreq > 0 ? e.memreq : (int)((unsigned)w.totmem[d] * (unsigned)e.mempct) / 100;
      fc.tmp_cores[fc.ntmp] = online ? 0 : e.coresreq;
      fc.tmp_aws[fc.ntmp] = (int)DF_MODE(w.dflags[d]);      // CustomInfo["QosPolicy"] cloned at Fit time (addJitteryQos compares against it)
      fc.ntmp++;
    }
    return true;
  }
  if (e.flags & EF_ENFLAME) {
    // enflame shared vGCU, enflame/device.go:382-478: the request becomes one profile of the node, then the gates
    int sel = -1;
    if (cnt > 0 && ka.mig_tab != nullptr) sel = hs_enf_select(ka.mig_tab[PK_GEOM(w.packed[td[0]])], e.memreq, e.mempct, e.coresreq);
    const MigCand* pc = sel >= 0 ? &ka.mig_tab[PK_GEOM(w.packed[td[0]])].cand[sel] : nullptr;
    if (pc == nullptr || pc->mem < 0) { REASON(14 /*ModeNotFit*/); return false; }
    const int pmem = pc->mem, pcore = (int)pc->core, slices = hs_enf_i32(*pc, 0);
    uint32_t useMaskE = 0xffu, nouseMaskE = 0u;
    if (e.has_use) useMaskE = uuid_mask_lookup(ka.uuid_pool, e.use_off, e.use_cnt, node);
    if (e.nouse_cnt) nouseMaskE = uuid_mask_lookup(ka.uuid_pool, e.nouse_off, e.nouse_cnt, node);
    for (int i = cnt - 1; i >= 0; i--) {
      const int d = td[i];
      if (!(w.dflags[d] & DF_HEALTH)) { REASON(6); continue; }
      if (!((e.type_allow >> PK_TYPE(w.packed[d])) & 1ull)) { REASON(0); continue; }
      if (!((useMaskE >> d) & 1u) || ((nouseMaskE >> d) & 1u)) { REASON(1); continue; }
      if (w.count[d] <= w.used[d]) { REASON(2); continue; }
      if (isMutex && w.used[d] > 0) { REASON(9); continue; }
      if (w.totmem[d] - w.mem[d] < pmem) { REASON(4); continue; }
      if (w.totcore[d] > 0 && w.totcore[d] - w.cores[d] < pcore) { REASON(5); continue; }
      if (k_nums > 0) {
        k_nums--;
        fc.tmp_dev[fc.ntmp] = d;
        fc.tmp_mem[fc.ntmp] = pmem;
        fc.tmp_cores[fc.ntmp] = pcore;
        fc.tmp_aws[fc.ntmp] = sel | (slices << 8);       // the profile and its slice count ride along to AddResourceUsage
        fc.ntmp++;
      }
      if (k_nums == 0) return true;
    }
    if (MODE == EV_EXPLAIN && fc.ntmp > 0) out.reasons[12] = (uint32_t)fc.ntmp;
    return false;
  }
  if (e.flags & EF_KUNLUN) {
    // kunlun whole XPUs: graghSelect (kunlun/topo.go:206-252) over the typeDevices in LIST order; fitFn = both uuid pairs
    // (merged into one use / no-use list on the host) and FitXPU = idle and healthy (device.go:196-206,238-240)
    uint32_t useMaskK = 0xffu, nouseMaskK = 0u;
    if (e.has_use) useMaskK = uuid_mask_lookup(ka.uuid_pool, e.use_off, e.use_cnt, node);
    if (e.nouse_cnt) nouseMaskK = uuid_mask_lookup(ka.uuid_pool, e.nouse_off, e.nouse_cnt, node);
    uint32_t fit = 0;              // by list position
    int idx[HS_D];
    int mism = 0, left = 0, right = 0, excl = 0;
    const bool vxpu = e.flags & EF_VXPU;
    for (int i = 0; i < cnt; i++) {
      const int d = td[i];
      idx[i] = (int)PK_INDEX(w.packed[d]);
      const bool uu = ((useMaskK >> d) & 1u) && !((nouseMaskK >> d) & 1u);
      // FitVXPU, vdevice.go:293-306 (int32 arithmetic as in Go)
      const bool vfit = (w.dflags[d] & DF_HEALTH) && !((int)((unsigned)e.memreq + (unsigned)w.mem[d]) > w.totmem[d]) &&
                        (w.used[d] == 0 || w.mem[d] / w.used[d] == e.memreq);
      if (vxpu && isMutex && w.used[d] > 0 && vfit) excl++;        // vdevice.go:259-266 counts these whatever the uuid lists say
      if (!uu) { mism++; continue; }
      const bool ok = vxpu ? ((!isMutex || w.used[d] == 0) && vfit) : (w.used[d] == 0 && (w.dflags[d] & DF_HEALTH));
      if (ok) { fit |= 1u << i; if (i < 4) left++; else right++; }
    }
    const int count = e.nums;
    int alloc[HS_D], na = 0;
    if (count == 8) {
      if (left + right == 8) for (int i = 0; i < 8; i++) alloc[na++] = i;
    } else if (count == 1 || count == 2 || count == 4) {
      int start = -1;
      if (left >= count || right >= count)
        for (int slots = count; slots <= 4 && start < 0; slots++) {
          const int num = (count & 1) ? (slots == 1 ? 1 : slots == 2 ? 3 : slots == 3 ? 2 : 4) : slots;     // oddorder {1,3,2,4}
          if (left == num) start = 0;
          else if (right == num) start = 4;
        }
      if (start >= 0) {                                           // devicepick, :192-204
        for (int t = start; t < 8 && t < cnt && na < count; t++) if ((fit >> t) & 1u) alloc[na++] = idx[t];
      } else if (count == 2) {                                    // interconnect, :112-131: the first inter-group pair in walk order
        for (int a = 0; a < cnt && na == 0; a++) {
          if (!((fit >> a) & 1u)) continue;
          for (int b = 0; b < cnt && na == 0; b++) {
            if (!((fit >> b) & 1u) || idx[b] == idx[a]) continue;
            const int lo = idx[a] < idx[b] ? idx[a] : idx[b], hi = idx[a] < idx[b] ? idx[b] : idx[a];
            if (lo < 4 && hi == lo + 4) { alloc[0] = idx[a]; alloc[1] = idx[b]; na = 2; }
          }
        }
      } else if (count == 4) {                                    // :132-160
        uint32_t unused = 0;
        int nun = 0;
        for (int i = 0; i < cnt; i++) if ((fit >> i) & 1u) { nun++; if (idx[i] < 32) unused |= 1u << idx[i]; }
        const uint32_t inter2[6] = {0x33u, 0xccu, 0x55u, 0xaau, 0x99u, 0x66u};   // 0-1-4-5, 2-3-6-7, 0-2-4-6, 1-3-5-7, 0-3-4-7, 1-2-5-6
        const uint32_t order[6][4] = {{0, 1, 4, 5}, {2, 3, 6, 7}, {0, 2, 4, 6}, {1, 3, 5, 7}, {0, 3, 4, 7}, {1, 2, 5, 6}};
        int chosen = -1;
        if (nun == 4 || nun == 5) {
          for (int c = 0; c < 6 && chosen < 0; c++) if ((unused & inter2[c]) == inter2[c]) chosen = c;
        } else if (nun == 6) {
          for (int c = 0; c < 6; c++) {
            if ((unused & inter2[c]) != inter2[c]) continue;
            chosen = c;
            const uint32_t rest = unused & ~inter2[c];             // delta: two cards; connected = an inter-group or in-group pair
            bool connected = false;
            for (int a = 0; a < 8; a++)
              for (int b = a + 1; b < 8; b++)
                if (((rest >> a) & 1u) && ((rest >> b) & 1u) && (b == a + 4 || (a < 4) == (b < 4))) connected = true;
            if (connected) break;
          }
        }
        if (chosen >= 0) for (int i = 0; i < 4; i++) alloc[na++] = (int)order[chosen][i];
      }
    }
    if (na == 0) {
      if (MODE == EV_EXPLAIN) {
        if (excl > 0) out.reasons[9] += (uint32_t)excl;
        else if (mism > 0) out.reasons[1] += (uint32_t)mism;
        else out.reasons[8]++;
      }
      return false;
    }
    for (int a = 0; a < na; a++)
      for (int i = 0; i < cnt; i++) {
        const int d = td[i];
        if ((int)PK_INDEX(w.packed[d]) != alloc[a]) continue;
        fc.tmp_dev[fc.ntmp] = d;
        fc.tmp_mem[fc.ntmp] = vxpu ? e.memreq : w.totmem[d];
        fc.tmp_cores[fc.ntmp] = vxpu ? e.coresreq : w.totcore[d];
        fc.ntmp++;
        break;
      }
    return true;
  }
  if (e.flags & EF_NEURON) {
    // AWS Neuron, awsneuron/device.go:443-541.  Totalcore / Usedcores are core bit masks; w.aws is the working copy's
    // CustomInfo["awsusageinfo"]; DF_NET carries CustomInfo["AWSNodeType"]: 255 absent, 1 = the instance type contains
    // "inf"/"Inf" (a ring), 0 = any other.
    const int require = e.coresreq;
    if (e.nums > 1) {
      // graphSelect (:406-441) over list POSITIONS, then the device whose Index equals each chosen position (:456-472)
      const int count = e.nums;
      int start_found = -1;
      const uint32_t nt = cnt > 0 ? DF_NET(w.dflags[td[0]]) : 255u;
      const bool ring = nt == 1u;
      if (nt != 255u && (ring || count == 4 || count == 8 || count == 16)) {
        for (int start = 0; start + count <= cnt && start_found < 0; start += ring ? 1 : count) {
          bool free_run = true;
          for (int it = start; it < start + count; it++)
            if (w.used[td[it]] > 0 || !(w.dflags[td[it]] & DF_HEALTH)) { free_run = false; break; }
          if (free_run) start_found = start;
        }
      }
      if (start_found < 0) { REASON(8 /*NumaNotFit*/); return false; }
      for (int pos = start_found; pos < start_found + count; pos++)
        for (int i = 0; i < cnt; i++) {
          const int d = td[i];
          if ((int)PK_INDEX(w.packed[d]) != pos) continue;
          const int prev = w.aws[d] < 0 ? 0 : w.aws[d];
          fc.tmp_dev[fc.ntmp] = d;
          fc.tmp_mem[fc.ntmp] = w.totmem[d];
          fc.tmp_cores[fc.ntmp] = w.totcore[d];
          fc.tmp_aws[fc.ntmp] = (prev == 0 && require == 2) ? 3 : ((prev == 0 && require == 1) ? 1 : 3 - prev);   // addCoreUsage, :367-389
          fc.ntmp++;
          break;
        }
      return true;
    }
    uint32_t useMaskN = 0xffu, nouseMaskN = 0u;
    if (e.has_use) useMaskN = uuid_mask_lookup(ka.uuid_pool, e.use_off, e.use_cnt, node);
    if (e.nouse_cnt) nouseMaskN = uuid_mask_lookup(ka.uuid_pool, e.nouse_off, e.nouse_cnt, node);
    for (int i = cnt - 1; i >= 0; i--) {
      const int d = td[i];
      if (w.aws[d] < 0) w.aws[d] = w.cores[d];        // :485-488, before every gate
      if (!(w.dflags[d] & DF_HEALTH)) { REASON(6); continue; }
      if (!((e.type_allow >> PK_TYPE(w.packed[d])) & 1ull)) { REASON(0); continue; }
      if (!((useMaskN >> d) & 1u) || ((nouseMaskN >> d) & 1u)) { REASON(1); continue; }
      if (w.count[d] <= w.used[d]) { REASON(2); continue; }
      if (isMutex && w.used[d] > 0) { REASON(9); continue; }
      const int have = (w.totcore[d] > 0 ? __popc((unsigned)w.totcore[d]) : 0) - (w.cores[d] > 0 ? __popc((unsigned)w.cores[d]) : 0);
      if (have < require) { REASON(5); continue; }     // countMaskAvailable, :357-365
      const int prev = w.aws[d];
      const int v = (prev == 0 && require == 2) ? 3 : ((prev == 0 && require == 1) ? 1 : 3 - prev);
      fc.tmp_dev[0] = d; fc.tmp_mem[0] = 0; fc.tmp_cores[0] = v; fc.tmp_aws[0] = v;
      fc.ntmp = 1;
      return true;
    }
    return false;
  }
  if (isAscend) {
    if (isHAMiCore && !(nflags & NF_VNPU_CORE)) {  // ascend/device.go:469-473
      REASON(14 /*ModeNotFit*/);
      return false;
    }
    if (e.flags & EF_ASCEND910) {                  // hasNetworkID, ascend/device.go:476-480
      bool all = true;
      for (int i = 0; i < cnt; i++) if (DF_NET(w.dflags[td[i]]) == 255u) { all = false; break; }
      if (all) needTopology = true;
    }
  }
  const bool is910CMulti = isAscend && (e.flags & EF_ASCEND910C) && origin > 1;
  const bool decrement = isNv ? !needTopology : (isAscend ? (!needTopology && !is910CMulti) : true);
  uint32_t useMask = 0xffu, nouseMask = 0u;
  if (e.has_use) useMask = uuid_mask_lookup(ka.uuid_pool, e.use_off, e.use_cnt, node);
  if (e.nouse_cnt) nouseMask = uuid_mask_lookup(ka.uuid_pool, e.nouse_off, e.nouse_cnt, node);
  const bool numaAssert = isNv && (pod.flags & PF_NUMA_ASSERT);
  // vastai die mode: isDieMode reads the FIRST device of the list (vastai/device.go:389-403); the AIC rides in NetworkID
  const bool vastaiDie = (e.flags & EF_VASTAI) && cnt > 0 && DF_NET(w.dflags[td[0]]) != 255u;
  if (vastaiDie) needTopology = true;       // every fitting die is collected, the choice goes by card (:306-320)

  for (int i = cnt - 1; i >= 0; i--) {
    const int d = td[i];
    const uint32_t df = w.dflags[d];
    const uint32_t pk = w.packed[d];
    if (!(df & DF_HEALTH)) { REASON(6); continue; }
    if (isNv && (df & DF_CORDON)) { REASON(7); continue; }
    bool found = (e.type_allow >> PK_TYPE(pk)) & 1ull;
    if (isNv && !((e.mode_allow >> DF_MODE(df)) & 1u)) found = false;
    if (!found) { REASON(0); continue; }
    if (e.flags & EF_WHOLE_CARD) {        // enflame/gcu.go:159-176: a card in use at all is a conflict; no uuid / slot gates
      if (w.used[d] > 0) { REASON(9); continue; }
      if (k_nums > 0) {
        k_nums--;
        fc.tmp_dev[fc.ntmp] = d;
        fc.tmp_mem[fc.ntmp] = e.memreq;
        fc.ntmp++;
      }
      if (k_nums == 0) return true;
      continue;
    }
    const int numa = PK_NUMA(pk);
    if (numaAssert && prevnuma != numa) {
      if (k_nums != origin) { if (MODE == EV_EXPLAIN) out.reasons[8] += (uint32_t)fc.ntmp; }
      k_nums = origin;
      prevnuma = numa;
      fc.ntmp = 0;
    }
    if (!((useMask >> d) & 1u) || ((nouseMask >> d) & 1u)) { REASON(1); continue; }
    if (w.count[d] <= w.used[d]) { REASON(2); continue; }
    if (isMutex && w.used[d] > 0) { REASON(9); continue; }
    if (countOnly) {                      // biren/device.go:226-241: no memory / core gates, the pick carries no memory
      if (vastaiDie && isMutex && DF_NET(df) != 255u) {
        // under mutex a card is exclusive: an idle die on a card with a die in use is a conflict (vastai/device.go:241-250,283-297)
        bool busy = false;
        for (int t = 0; t < cnt; t++) if (w.used[td[t]] > 0 && DF_NET(w.dflags[td[t]]) == DF_NET(df)) busy = true;
        if (busy) { REASON(9); continue; }
      }
      if (k_nums > 0) {
        if (!vastaiDie) k_nums--;
        fc.tmp_dev[fc.ntmp] = d;
        fc.tmp_mem[fc.ntmp] = 0;
        fc.ntmp++;
      }
      if (k_nums == 0 && !vastaiDie) return true;
      continue;
    }
    if (e.flags & EF_AMD) {               // amd/device.go:308-350
      int memReq = e.memreq;
      if (memReq <= 0 && w.totmem[d] > 0) memReq = w.totmem[d];
      if (w.totmem[d] - w.mem[d] < memReq) { REASON(4); continue; }
      int coreReq = 0;
      if (e.coresreq > 0) {
        if (w.totcore[d] <= 0) { REASON(5); continue; }
        coreReq = (int)((unsigned)w.totcore[d] * (unsigned)e.coresreq) / 100;
        coreReq = coreReq < 1 ? 1 : coreReq;
        coreReq = coreReq > w.totcore[d] ? w.totcore[d] : coreReq;
      } else if (w.totmem[d] > 0 && memReq >= w.totmem[d]) {
        coreReq = w.totcore[d];
      }
      if (w.totcore[d] - w.cores[d] < coreReq) { REASON(5); continue; }
      if (k_nums > 0) {
        k_nums--;
        fc.tmp_dev[fc.ntmp] = d;
        fc.tmp_mem[fc.ntmp] = memReq;
        fc.tmp_cores[fc.ntmp] = coreReq;
        fc.ntmp++;
      }
      if (k_nums == 0) return true;
      continue;
    }
    int memreq = 0;
    if (e.memreq > 0) memreq = e.memreq;
    if (e.mempct != 101 && e.memreq == 0) memreq = (int)((unsigned)w.totmem[d] * (unsigned)e.mempct) / 100;
    // fitQuota, nvidia/device.go:784-788 and cambricon/device.go:439-443
    if (quota_on && !quota_fits(*qa, qv, fc.tmp_dev, fc.tmp_mem, fc.ntmp, d, memreq, fc.cores)) { REASON(13); continue; }
    if (w.totmem[d] - w.mem[d] < memreq) { REASON(4); continue; }
    const int etc = isHAMiCore ? 100 : w.totcore[d];
    if (etc - w.cores[d] < fc.cores) { REASON(5); continue; }
    if (etc == 100 && fc.cores == 100 && w.used[d] > 0) { REASON(9); continue; }
    if (etc != 0 && w.cores[d] == etc && fc.cores == 0) { REASON(3); continue; }
    const bool migDev = isNv && DF_MODE(df) == DF_MODE_MIG;
    if (migDev) {
      // CustomFilterRule, nvidia/device.go:602-619 with the resolved memory (:811-814): the picks this Fit already queued
      // on the same GPU take their placements first, then the request must still find a profile with a free placement
      bool ok = ka.mig_tab != nullptr;
      if (ok) {
        const MigGeom& g = ka.mig_tab[PK_GEOM(pk)];
        uint32_t occ = DF_OCC(df);
        for (int t = 0; t < fc.ntmp && ok; t++) {
          if (fc.tmp_dev[t] != d) continue;
          const int sel = hs_mig_select(g, occ, fc.tmp_mem[t]);
          if (sel < 0) ok = false; else occ |= (uint32_t)sel >> 8;
        }
        if (ok) ok = hs_mig_select(g, occ, memreq) >= 0;
      }
      if (!ok) { REASON(11 /*CardMigTopologyInfeasible*/); continue; }
    }
    // customFilterRule, mthreads/device.go:266-276,362-366: the pod's earlier containers all sit on this very card
    if ((e.flags & EF_MTHREADS) && (w.mth_prev & ~(1u << d))) { REASON(10); continue; }
    if (k_nums > 0) {
      if (decrement) k_nums--;
      fc.tmp_dev[fc.ntmp] = d;
      fc.tmp_mem[fc.ntmp] = memreq;
      fc.ntmp++;
    }
    if (k_nums == 0 && decrement) return true;
    if (migDev && fc.ntmp < HS_D) i++;      // nvidia/device.go:846-848: a GPU in MIG mode is visited again for the next instance
  }

  if (is910CMulti) {
    // computeBestCombination910C, ascend/device.go:677-733 (equal-size cards in ascending card id)
    int sel_dev[HS_D], sel_mem[HS_D], nsel = 0;
    int task = origin;
    // indexToDevice is keyed by Idx: a later candidate with the same Idx replaces an earlier one
    for (int pass = 0; pass < 1; pass++) {
      // full cards (both NPUs present among the candidates), ascending card id
      for (int card = 0; card < 128 && task > 0; card++) {
        int members[HS_D], nm = 0;
        for (int t = 0; t < fc.ntmp; t++)
          if (PK_INDEX(w.packed[fc.tmp_dev[t]]) / 2 == card) members[nm++] = t;
        if (nm == 0) continue;
        if (nm == 2) {
          for (int m = 0; m < nm && nsel < HS_D; m++) {
            // map the index back through indexToDevice (last candidate with that Idx wins)
            const int idx = PK_INDEX(w.packed[fc.tmp_dev[members[m]]]);
            int pick = members[m];
            for (int t = 0; t < fc.ntmp; t++) if (PK_INDEX(w.packed[fc.tmp_dev[t]]) == idx) pick = t;
            sel_dev[nsel] = fc.tmp_dev[pick];
            sel_mem[nsel] = fc.tmp_mem[pick];
            nsel++;
          }
          task -= 2;
        }
      }
    }
    if (nsel != origin) {
      if (MODE == EV_EXPLAIN) {
        out.reasons[12] = (uint32_t)nsel; out.fcards = (uint32_t)origin;
        if (nsel == 0) out.fkind = FK_HIST_ZALLOC;      // the reference still lists "0/<request> AllocatedCardsInsufficientRequest"
      }
      return false;
    }
    for (int t = 0; t < nsel; t++) { fc.tmp_dev[t] = sel_dev[t]; fc.tmp_mem[t] = sel_mem[t]; }
    fc.ntmp = nsel;
    return true;
  }

  if (needTopology) {
    if (fc.ntmp == origin) return true;
    if (fc.ntmp > origin) {
      const int m = fc.ntmp;
      if (isNv) {
        const bool uniform = nflags & NF_PAIR_UNIFORM;
        const int16_t* P = ka.pair + (size_t)node * 64;
        if (origin == 1) {
          // computeWorstSingleCard, nvidia/device.go:927-951: first strictly-lowest Σ pair score
          int best = 0;
          if (!uniform) {
            int worst = 0;
            bool have = false;
            for (int a = 0; a < m; a++) {
              int total = 0;
              for (int b = 0; b < m; b++) if (b != a) total += (int)P[fc.tmp_dev[a] * 8 + fc.tmp_dev[b]];
              if (!have || total < worst) { have = true; worst = total; best = a; }
            }
          }
          fc.tmp_dev[0] = fc.tmp_dev[best];
          fc.tmp_mem[0] = fc.tmp_mem[best];
          fc.ntmp = 1;
          return true;
        }
        // generateCombinations + computeBestCombination, nvidia/device.go:887-915,953-978:
        // lexicographic enumeration, first strictly-highest Σ pairwise score wins.
        uint32_t bestMask = (1u << origin) - 1u;  // the first combination
        if (!uniform) {
          int bestScore = 0;
          bool have = false;
          for (uint32_t mask = 1; mask < (1u << m); mask++) {
            if (__popc(mask) != origin) continue;
            int total = 0;
            for (int a = 0; a < m; a++) {
              if (!((mask >> a) & 1u)) continue;
              for (int b = a + 1; b < m; b++)
                if ((mask >> b) & 1u) total += (int)P[fc.tmp_dev[a] * 8 + fc.tmp_dev[b]];
            }
            bool take;
            if (!have) take = true;
            else if (total != bestScore) take = total > bestScore;
            else {  // equal score: the lexicographically earlier combination was enumerated first
              const uint32_t diff = mask ^ bestMask;
              take = (mask & (diff & (0u - diff))) != 0;
            }
            if (take) { have = true; bestScore = total; bestMask = mask; }
          }
        }
        int n = 0;
        for (int a = 0; a < m; a++)
          if ((bestMask >> a) & 1u) { fc.tmp_dev[n] = fc.tmp_dev[a]; fc.tmp_mem[n] = fc.tmp_mem[a]; n++; }
        fc.ntmp = n;
        return true;
      }
      // Ascend NetworkID topology, ascend/device.go:599-607,631-675
      if (origin == 1) { fc.ntmp = 1; return true; }
      int sel_dev[HS_D], sel_mem[HS_D], nsel = 0;
      uint32_t taken = 0;  // candidates already emitted
      while (nsel < origin) {
        // next group: largest remaining count, ties → ascending NetworkID (canonical rule 7)
        int bestNet = -1, bestCnt = 0;
        for (int a = 0; a < m; a++) {
          if ((taken >> a) & 1u) continue;
          const int net = (int)DF_NET(w.dflags[fc.tmp_dev[a]]);
          int c = 0;
          for (int b = 0; b < m; b++)
            if (!((taken >> b) & 1u) && (int)DF_NET(w.dflags[fc.tmp_dev[b]]) == net) c++;
          if (c > bestCnt || (c == bestCnt && net < bestNet)) { bestCnt = c; bestNet = net; }
        }
        if (bestNet < 0) break;
        for (int a = 0; a < m && nsel < origin; a++) {
          if ((taken >> a) & 1u) continue;
          if ((int)DF_NET(w.dflags[fc.tmp_dev[a]]) != bestNet) continue;
          sel_dev[nsel] = fc.tmp_dev[a];
          sel_mem[nsel] = fc.tmp_mem[a];
          nsel++;
        }
        for (int a = 0; a < m; a++)
          if ((int)DF_NET(w.dflags[fc.tmp_dev[a]]) == bestNet) taken |= 1u << a;
      }
      for (int t = 0; t < nsel; t++) { fc.tmp_dev[t] = sel_dev[t]; fc.tmp_mem[t] = sel_mem[t]; }
      fc.ntmp = nsel;
      return true;
    }
  }
  if (fc.ntmp > 0) { if (MODE == EV_EXPLAIN) out.reasons[12] = (uint32_t)fc.ntmp; }
  return false;
}

// fitInDevices, score.go:81-137, for the entries [eb, ee) of one container
template <int MODE>
__device__ bool dev_fit_container(NodeWork& w, const DevPod& pod, const DevEntry* ent, int eb, int ee,
                                  uint32_t node, uint32_t nflags, const KernelArgs& ka, EvalOut& out,
                                  float* score_acc, uint32_t* score_zero, QuotaAcc* qa) {
  for (int d = 0; d < w.ndev; d++) w.score[d] = dev_compute_score(w, d, pod, ent, eb, ee);
  for (int ei = eb; ei < ee; ei++) {
    const DevEntry& e = ent[ei];
    dev_sort(w, pod);
    if (e.vendor == HV_NONE) {
      if (MODE == EV_EXPLAIN) out.fkind = FK_TYPE_NOT_FOUND;
      return false;
    }
    int td[HS_D], cnt = 0;
    bool anyMig = false;
    for (int i = 0; i < w.ndev; i++) {
      const int d = ord_get(w.order, i);
      if ((e.type_contains >> PK_TYPE(w.packed[d])) & 1ull) {
        td[cnt++] = d;
        if (DF_MODE(w.dflags[d]) == DF_MODE_MIG) anyMig = true;
      }
    }
    // isMIGRequest, score.go:54-69
    const bool migReq = e.vendor == HV_NVIDIA && (e.flags & EF_MIGMODE_OK) && anyMig;
    if (e.nums > cnt && !migReq) {
      if (MODE == EV_EXPLAIN) out.fkind = FK_NODE_INSUFFICIENT;
      return false;
    }
    FitCtx fc;
    if (!dev_fit<MODE>(w, pod, e, td, cnt, node, nflags, ka, fc, out, qa)) return false;
    if (qa && (e.vendor == HV_NVIDIA || e.vendor == HV_MLU))
      quota_append(*qa, e.vendor == HV_NVIDIA ? 0 : 1, fc.tmp_dev, fc.tmp_mem, fc.ntmp, fc.cores);
    // AddResourceUsage per pick (nvidia/device.go:645-664 non-MIG; identical for the other vendors)
    if (e.flags & EF_MTHREADS)
      for (int t = 0; t < fc.ntmp; t++) w.mth_prev |= 1u << fc.tmp_dev[t];   // devinput[k.Type] grows by this container (score.go:132)
    int netcnt_max = 0;
    bool missing_net = false;
    for (int t = 0; t < fc.ntmp; t++) {
      const int d = fc.tmp_dev[t];
      int pick_cores = (e.flags & (EF_AMD | EF_NEURON)) ? fc.tmp_cores[t] : fc.cores;
      int mig_sel = -1;
      if (e.vendor == HV_NVIDIA && DF_MODE(w.dflags[d]) == DF_MODE_MIG) {
        // AddResourceUsage on a GPU in MIG mode (nvidia/device.go:645-659): the pick becomes the chosen profile
        mig_sel = ka.mig_tab != nullptr ? hs_mig_select(ka.mig_tab[PK_GEOM(w.packed[d])], DF_OCC(w.dflags[d]), fc.tmp_mem[t]) : -1;
        if (mig_sel < 0) {
          if (MODE == EV_EXPLAIN) { out.fkind = FK_MIG_ADD_FAILED; out.fcards = (uint32_t)d; }
          return false;
        }
        const MigCand& mc = ka.mig_tab[PK_GEOM(w.packed[d])].cand[mig_sel & 15];
        fc.tmp_mem[t] = mc.mem;
        pick_cores = mc.core;
        w.dflags[d] |= ((uint32_t)mig_sel >> 8) << 16;      // MigAllocationsInUse of the working copy
      }
      if (e.flags & EF_NEURON) {          // awsneuron/device.go:340-355
        if (w.aws[d] < 0) w.aws[d] = 0;
        w.aws[d] += fc.tmp_aws[t];
      }
      if (e.flags & EF_SGPU) {            // metax/sdevice.go:310-330
        pick_cores = fc.tmp_cores[t];
        mig_sel = fc.tmp_aws[t] & 7;      // the device's policy at Fit time rides in the row's spare bits (AL_MIG_CAND)
        w.used[d] += 1;
        w.cores[d] += pick_cores;
        w.mem[d] += fc.tmp_mem[t];
        if ((pod.sg_class & 3) != 1) {    // not an online pod: the working copy's policy becomes the pod's ("" for a whole card)
          const uint32_t q = pick_cores == 100 ? 0u : (uint32_t)pod.sg_qos;
          w.dflags[d] = (w.dflags[d] & ~(7u << 3)) | ((q & 7u) << 3);
        }
        w.sg_picks |= 1u << d;
        if (pick_cores == 100 && ((w.sg_excl >> (4 * d)) & 15u) < 15u) w.sg_excl += 1u << (4 * d);   // a card an init AND an app container took counts twice
      } else if (e.flags & EF_ENFLAME) {         // enflame/device.go:372-380: Used grows by the profile's slice count
        pick_cores = fc.tmp_cores[t];
        mig_sel = fc.tmp_aws[t] & 15;     // reported like a MIG pick: the candidate index in the row's spare bits
        int s = fc.tmp_aws[t] >> 8;
        s = s <= 0 ? 1 : s;
        const long long u = (long long)w.used[d] + s;
        w.used[d] = u > 2147483647LL ? 2147483647 : (int)u;
        w.cores[d] += pick_cores;
        w.mem[d] += fc.tmp_mem[t];
      } else if (e.flags & EF_VXPU) {
        pick_cores = fc.tmp_cores[t];     // the request's Coresreq as asked (vdevice.go:283)
        w.used[d] += 1;
        w.cores[d] += pick_cores;
        w.mem[d] += fc.tmp_mem[t];
      } else if (e.flags & EF_KUNLUN) {   // kunlun/device.go:179-182: AddResourceUsage counts the slot only; the pick keeps the whole card
        pick_cores = fc.tmp_cores[t];
        w.used[d] += 1;
        const int ki = (int)PK_INDEX(w.packed[d]);
        if (ki < 32) w.kl_cur |= 1u << ki;
      } else {
        w.used[d] += 1;
        w.cores[d] += pick_cores;
        w.mem[d] += fc.tmp_mem[t];
      }
      if (MODE == EV_RECORD && out.n_alloc < HS_MAX_ALLOC) {
        DevAlloc a;
        a.entry = (uint8_t)(ei);
        a.dev = (uint8_t)d;
        if (mig_sel >= 0) { a.entry |= (uint8_t)((mig_sel & 15) << 4); a.dev |= (uint8_t)(8u | (((uint32_t)mig_sel >> 4) & 7u) << 4); }
        a.usedcores = (int16_t)pick_cores;
        a.usedmem = fc.tmp_mem[t];
        out.allocs[out.n_alloc++] = a;
      }
      if (e.score_slot >= 0 && !(e.flags & EF_METAX)) {
        const int net = (int)DF_NET(w.dflags[d]);
        if (net == 255) missing_net = true;
        int c = 0;
        for (int u = 0; u < fc.ntmp; u++) if ((int)DF_NET(w.dflags[fc.tmp_dev[u]]) == net) c++;
        if (c > netcnt_max) netcnt_max = c;
      }
    }
    // Ascend ScoreNode per container, ascend/device.go:388-424
    if (e.score_slot >= 0 && (e.flags & EF_METAX)) {
      // metax ScoreNode (metax/device.go:185-214) only counts the pod's devices: the slot's accumulator is that count
      score_acc[e.score_slot] = __fadd_rn(score_acc[e.score_slot], (float)fc.ntmp);
    } else if (e.score_slot >= 0 && fc.ntmp > 0) {
      if (missing_net) *score_zero |= 1u << e.score_slot;
      else score_acc[e.score_slot] = __fadd_rn(score_acc[e.score_slot], __fdiv_rn((float)netcnt_max, (float)fc.ntmp));
    }
  }
  return true;
}

// scoreNode, score.go:294-346.  `base` usage must already be in w.used/mem/cores.
// Returns fit; *score_out = node score (pre-allocation default score + vendor ScoreNode terms).
template <int MODE>
__device__ bool dev_score_node(NodeWork& w, const DevPod& pod, const DevEntry* ent, uint32_t node, uint32_t nflags,
                               const KernelArgs& ka, EvalOut& out, float* score_out) {
  float score_acc[HS_SCORE_SLOTS];
  uint32_t score_zero = 0;
#pragma unroll
  for (int s = 0; s < HS_SCORE_SLOTS; s++) score_acc[s] = 0.0f;
  const uint32_t identity = 0x76543210u;
  const float pre = node_default_score(w);  // on the pre-allocation usage (score.go:320-328)
  QuotaAcc qacc;
  QuotaAcc* qa = nullptr;
  if (ka.quota != nullptr && (pod.q_idx[0] >= 0 || pod.q_idx[1] >= 0)) { quota_load(qacc, pod, ka); qa = &qacc; }
  const bool base_key[2] = {(nflags & NF_BASE_NV) != 0, (nflags & NF_BASE_MLU) != 0};
  w.mth_prev = 0;
  w.kl_cur = 0;
  w.sg_picks = 0; w.sg_excl = 0;
  uint32_t sg_idle = 0;               // metax sGPU ScoreNode's `previous`: the devices idle before the pod (base usage)
  int sg_slot = -1;
  for (int e = 0; e < pod.n_entries; e++) if (ent[e].flags & EF_SGPU) sg_slot = ent[e].score_slot;
  if (sg_slot >= 0) for (int d = 0; d < w.ndev; d++) if (w.used[d] == 0) sg_idle |= 1u << d;
  uint32_t kl_prev = 0;               // kunlun ScoreNode's `previous`: cards in use before the pod (score.SnapshotDevice, score.go:328)
  bool has_kunlun = false;
  for (int e = 0; e < pod.n_entries; e++) if ((ent[e].flags & EF_KUNLUN) && !(ent[e].flags & EF_VXPU)) has_kunlun = true;
  if (has_kunlun)
    for (int d = 0; d < w.ndev; d++) if (w.used[d] > 0 && PK_INDEX(w.packed[d]) < 32) kl_prev |= 1u << PK_INDEX(w.packed[d]);
  uint32_t metax_slots = 0;           // score slots that belong to a metax type (their accumulator is a device count)
  uint32_t plain_slots = 0;           // … to a vastai type: ScoreNode is the plain sum, no weight (vastai/device.go:181-216)
  for (int e = 0; e < pod.n_entries; e++) {
    if ((ent[e].flags & EF_METAX) && ent[e].score_slot >= 0) metax_slots |= 1u << ent[e].score_slot;
    if ((ent[e].flags & (EF_VASTAI | EF_SGPU)) && ent[e].score_slot >= 0) plain_slots |= 1u << ent[e].score_slot;
  }

  if (pod.n_init > 0) {
    int bu[HS_D], bm[HS_D], bc[HS_D];
    uint32_t bf[HS_D];       // dflags carry the working copy's MIG reservations
    for (int d = 0; d < w.ndev; d++) { bu[d] = w.used[d]; bm[d] = w.mem[d]; bc[d] = w.cores[d]; bf[d] = w.dflags[d]; }
    for (int c = 0; c < pod.n_init && c < pod.n_ctr; c++) {
      const int eb = pod.ctr_begin[c], ee = pod.ctr_begin[c + 1];
      if (eb == ee) {          // len(req) == 0: every base type gets an empty container (score.go:233-238; the rows are padded on the host)
        if (qa) for (int v = 0; v < 2; v++) if (base_key[v]) qa->len[v]++;
        continue;
      }
      w.order = identity;      // node.DeepCopy(): registration order, fresh usage (score.go:241)
      for (int d = 0; d < HS_D; d++) w.aws[d] = -1;
      if (!dev_fit_container<MODE>(w, pod, ent, eb, ee, node, nflags, ka, out, score_acc, &score_zero, qa)) return false;
      if (qa) for (int v = 0; v < 2; v++) if (base_key[v] && qa->len[v] == c) qa->len[v]++;   // score.go:246-250
      for (int d = 0; d < w.ndev; d++) { w.used[d] = bu[d]; w.mem[d] = bm[d]; w.cores[d] = bc[d]; w.dflags[d] = bf[d]; }
    }
  }
  w.order = identity;
  w.mth_prev = 0;
  for (int d = 0; d < HS_D; d++) w.aws[d] = -1;   // appNodeCopy := node.DeepCopy(): no awsusageinfo key yet (scheduler.go:684 clones CustomInfo)
  if (qa) quota_reset_lists(*qa);     // score.Devices starts empty for the app containers (score.go:321-326)
  int app_index = 0;
  for (int c = pod.n_init; c < pod.n_ctr; c++) {
    const int eb = pod.ctr_begin[c], ee = pod.ctr_begin[c + 1];
    int sums = 0;
    for (int e = eb; e < ee; e++) sums += ent[e].nums;
    if (sums == 0) {
      if (qa) for (int v = 0; v < 2; v++) if (base_key[v]) qa->len[v]++;                         // score.go:266-271
      app_index++;
      continue;
    }
    if (!dev_fit_container<MODE>(w, pod, ent, eb, ee, node, nflags, ka, out, score_acc, &score_zero, qa)) return false;
    if (qa) for (int v = 0; v < 2; v++) if (base_key[v] && qa->len[v] == app_index) qa->len[v]++;  // score.go:278-282
    app_index++;
  }
  if (sg_slot >= 0 && w.sg_picks != 0u) {
    // MetaxSDevices.ScoreNode (metax/sdevice.go:295-304), then OverrideScore's +-10000 (policy/node_policy.go:82-88)
    int result = 0;
    if ((pod.sg_class & 3) == 1) {                              // scoreOnlineDevices, :772-811
      for (int d = 0; d < w.ndev; d++) {
        if (!((w.sg_picks >> d) & 1u)) continue;
        const int on = (int)SG_ON(w.dflags[d]), off = (int)SG_OFF(w.dflags[d]);
        result += (off > 0 && on == 0) ? -100 * off : (-20 * on - 2 * off);
      }
    } else if (w.sg_excl != 0u && (pod.sg_class & 4)) {          // topologyAwareEnable :647-667, scoreExclusiveDevices :669-
      auto zone = [&](int d) { const uint32_t df = w.dflags[d]; const int z = (df & DF_SGPU) ? (int)DF_NET(df) : 0; return z == 255 ? 0 : z; };
      uint32_t exmask = 0;
      for (int d = 0; d < w.ndev; d++) if ((w.sg_excl >> (4 * d)) & 15u) exmask |= 1u << d;
      auto pairs = [&](uint32_t mask, bool counted) {            // LinkDevices.Score, metax/score.go:27-57 (equal uuids score 0)
        int s = 0;
        for (int a = 0; a < w.ndev; a++)
          for (int b = a + 1; b < w.ndev; b++)
            if (((mask >> a) & 1u) && ((mask >> b) & 1u) && zone(a) != 0 && zone(a) == zone(b))
              s += counted ? 10 * (int)((w.sg_excl >> (4 * a)) & 15u) * (int)((w.sg_excl >> (4 * b)) & 15u) : 10;
        return s;
      };
      const int availableScore = pairs(sg_idle, false), allocatedScore = pairs(exmask, true), restScore = pairs(sg_idle & ~exmask, false);
      result = 10 * allocatedScore - (availableScore - allocatedScore - restScore);
    }
    score_acc[sg_slot] = __fmul_rn((pod.flags & PF_NODE_SPREAD) ? -10000.0f : 10000.0f, (float)result);
  }
  // OverrideScore, policy/node_policy.go:67-95: Σ vendor ScoreNode in sorted type-key order
  float devScore = 0.0f;
#pragma unroll
  for (int s = 0; s < HS_SCORE_SLOTS; s++) {
    float v = ((score_zero >> s) & 1u) ? 0.0f : (((plain_slots >> s) & 1u) ? score_acc[s] : __fmul_rn(score_acc[s], 10.0f));
    if ((metax_slots >> s) & 1u) {
      // 2000 - losses[n] under node policy "binpack", 2000 - scores[n] under "spread", 0 without the annotation
      const int cnt = (int)score_acc[s];
      const int16_t* T = ka.pair + (size_t)node * 64;
      v = 0.0f;
      if ((pod.flags & PF_NODE_BINPACK) && (nflags & NF_METAX_LOSS)) v = __fsub_rn(2000.0f, (float)(cnt < 32 ? (int)T[cnt] : 0));
      else if ((pod.flags & PF_NODE_SPREAD) && (nflags & NF_METAX_SCORE)) v = __fsub_rn(2000.0f, (float)(cnt < 32 ? (int)T[32 + cnt] : 0));
    }
    devScore = __fadd_rn(devScore, v);
  }
  if (has_kunlun && (nflags & NF_ALL_KUNLUN)) {
    // kunlun ScoreNode (device.go:154-177) + calcscore / countbubble (topo.go:50-95).  It is the only non-zero term on such
    // a node (every device is a kunlun card), so its place in the sum does not matter.
    int bubble[2];
    const uint32_t sets[2] = {kl_prev, w.kl_cur};
    for (int k = 0; k < 2; k++) {
      const int left = __popc(sets[k] & 0xfu), right = __popc(sets[k] >> 4);
      const int gl = left == 4 ? 0 : (left == 1 ? 2 : 1), gr = right == 4 ? 0 : (right == 1 ? 2 : 1);
      bubble[k] = (left == 0 && right == 0) ? 1 : gl + gr;
    }
    const int diff = bubble[1] - bubble[0];
    const float ks = diff == -1 ? 3000.0f : (diff == 0 ? 2000.0f : (diff == 2 ? 0.0f : 1000.0f));
    devScore = __fadd_rn(devScore, ks);
  }
  *score_out = __fadd_rn(pre, devScore);
  return true;
}

// Order-preserving key: best node == max key.  Bits 63-32: float32 score mapped to an unsigned that
// sorts like the float (inverted for node policy "spread", which picks the MIN score,
// node_policy.go:48-54); bits 31-8: candidate position + 1, so among equal scores the highest position
// wins (canonical rule 3) and 0 means "no node"; bits 7-0: reserved for the barrier epoch tag.
#define KEY_POS(k) ((int)(((k) >> 8) & 0xffffffull) - 1)
#define KEY_VALID(k) ((((k) >> 8) & 0xffffffull) != 0)
__device__ __forceinline__ unsigned long long make_key(float score, bool spread, uint32_t pos) {
  if (score == 0.0f) score = 0.0f;  // -0 == +0 in Go's comparison
  uint32_t u = __float_as_uint(score);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  if (spread) u = ~u;
  return ((unsigned long long)u << 32) | ((unsigned long long)((pos + 1u) & 0xffffffu) << 8);
}
