// hs_requests.cc — request extraction: device.Resourcereqs for one container
// (pkg/device/devices.go:639-691) over the vendors' GenerateResourceRequests:
//   NVIDIA  pkg/device/nvidia/device.go:520-600     DCU     pkg/device/hygon/device.go:168-239
//   MLU     pkg/device/cambricon/device.go:250-305  Ascend  pkg/device/ascend/device.go:303-378
//   and the vendors of hs_config.vendors: iluvatar, biren / vastai, amd, mthreads, enflame GCU (cited at each case)
// Input = the container's resource limits / requests with the three views of a resource.Quantity the reference
// uses (AsInt64, Value, Sign), output = the ContainerDeviceRequest list hs_pod carries.  Plain host code.
#include <stdint.h>
#include <string.h>

#include "../../include/hami_sched.h"

namespace {

const int64_t kMaxInt32 = 2147483647;

const hs_quantity* lookup(const hs_container_resources* c, const char* name) {
  if (!name || !name[0]) return nullptr;
  for (uint32_t i = 0; i < c->n_limits; i++)                  // Limits first, then Requests
    if (c->limits[i].name && strcmp(c->limits[i].name, name) == 0) return &c->limits[i];
  for (uint32_t i = 0; i < c->n_requests; i++)
    if (c->requests[i].name && strcmp(c->requests[i].name, name) == 0) return &c->requests[i];
  return nullptr;
}

const hs_quantity* lookup_limits(const hs_container_resources* c, const char* name) {
  if (!name || !name[0]) return nullptr;
  for (uint32_t i = 0; i < c->n_limits; i++)
    if (c->limits[i].name && strcmp(c->limits[i].name, name) == 0) return &c->limits[i];
  return nullptr;
}

// returns false for the zero-value ContainerDeviceRequest{}
bool generate(const hs_vendor_resources& v, const hs_container_resources* c, hs_container_request& r) {
  if (v.kind == HS_REQ_METAX_SGPU) {
    // metax/sdevice.go:222-293: Limits only; vcore defaults to 100; vmemory in Gi unless the quantity carries a unit
    r.key = v.common_word; r.type = v.common_word;
    const hs_quantity* q = lookup_limits(c, v.count_name);
    if (!q || !q->as_int64_ok) return false;
    if (q->as_int64 <= 0 || q->as_int64 > kMaxInt32) return false;
    int64_t core = 100, mem = 0;
    const hs_quantity* cq = lookup_limits(c, v.cores_name);
    if (cq && cq->as_int64_ok) {
      if (cq->as_int64 < 0 || cq->as_int64 > kMaxInt32) return false;
      core = cq->as_int64;
    }
    const hs_quantity* mq = lookup_limits(c, v.memory_name);
    if (mq && mq->as_int64_ok) {
      if (mq->has_suffix) mem = mq->as_int64 / 1024 / 1024;
      else {
        if (mq->as_int64 < 0 || mq->as_int64 > kMaxInt32 / 1024) return false;
        mem = mq->as_int64 * 1024;
      }
      if (mem < 0 || mem > kMaxInt32) return false;
    }
    r.nums = (int32_t)q->as_int64; r.memreq = (int32_t)mem; r.mempct = mem == 0 ? 100 : 0; r.coresreq = (int32_t)core;
    return true;
  }
  if (v.kind == HS_REQ_ENFLAME) {
    // enflame/device.go:322-366: a positive enflame.com/drs-gcu count is a DIRECT slice request; otherwise memory / core by
    // spec (getContainerResourceRequest :747-758: Quantity.Value() of Limits, then Requests)
    r.key = v.common_word; r.type = v.common_word; r.nums = 1;
    const hs_quantity* q = lookup(c, v.count_name);
    if (q && q->as_int64_ok && q->as_int64 > 0) {
      if (q->as_int64 > kMaxInt32) return false;
      r.memreq = (int32_t)q->as_int64; r.mempct = 0; r.coresreq = 0;
      return true;
    }
    const hs_quantity* mem = lookup(c, v.memory_name);
    const hs_quantity* core = lookup(c, v.cores_name);
    if (!mem && !core) return false;
    if (mem && mem->value > kMaxInt32) return false;
    if (core && core->value > kMaxInt32) return false;
    r.memreq = mem ? (int32_t)mem->value : 0;
    r.coresreq = core ? (int32_t)core->value : 0;
    r.mempct = -1;
    return true;
  }
  if (v.kind == HS_REQ_NEURON) {
    // awsneuron/device.go:287-334: by device (count_name) or, when that is absent, by NeuronCore (cores_name) through
    // splitCoreRequest (:141-154); validateResourceRequest (:115-129): an integer in (0, max]; default_cores carries the
    // vendor object's coresPerAWSNeuron (0 until a node registered)
    r.key = v.common_word; r.type = v.common_word; r.memreq = 0; r.mempct = 0;
    const hs_quantity* q = lookup(c, v.count_name);
    if (q) {
      if (!q->as_int64_ok || q->as_int64 <= 0 || q->as_int64 > kMaxInt32) return false;
      r.nums = (int32_t)q->as_int64;
      r.coresreq = v.default_cores;
      return true;
    }
    q = lookup(c, v.cores_name);
    if (!q) return false;
    if (!q->as_int64_ok || q->as_int64 <= 0 || q->as_int64 > kMaxInt32 * 2) return false;
    const int64_t cores = q->as_int64;
    const int64_t per = v.default_cores <= 0 ? 2 : (v.default_cores < 2 ? v.default_cores : 2);   // coresPerDevice, :131-137
    if (cores <= per) { r.nums = 1; r.coresreq = (int32_t)cores; return true; }
    if (cores % per != 0) return false;
    if (cores / per > kMaxInt32) return false;
    r.nums = (int32_t)(cores / per);
    r.coresreq = (int32_t)per;
    return true;
  }
  const hs_quantity* cnt = lookup(c, v.count_name);
  if (!cnt || !cnt->as_int64_ok) return false;
  const int64_t n = cnt->as_int64;
  const int64_t factor = v.memory_factor > 1 ? v.memory_factor : 1;
  const hs_quantity* mem = lookup(c, v.memory_name);
  const hs_quantity* core = lookup(c, v.cores_name);
  int64_t memnum = 0;
  int32_t mempnum, corenum;
  switch (v.kind) {
    case HS_REQ_NVIDIA: {
      if (mem) {
        if (!mem->as_int64_ok || mem->as_int64 < 0 || mem->as_int64 > kMaxInt32 / factor) return false;
        memnum = mem->as_int64 * factor;
      }
      mempnum = 101;                                          // "unset" sentinel (nvidia/device.go:551)
      const hs_quantity* pct = lookup(c, v.memory_percentage_name);
      if (pct && pct->as_int64_ok) {
        int64_t p = pct->as_int64 > 100 ? 100 : pct->as_int64;
        mempnum = p > 0 ? (int32_t)p : 101;
      }
      if (mempnum == 101 && memnum == 0) {
        if (v.default_memory != 0) memnum = v.default_memory;
        else mempnum = 100;
      }
      corenum = v.default_cores;
      if (core && core->as_int64_ok) corenum = (int32_t)core->as_int64;
      break;
    }
    case HS_REQ_DCU: {
      if (n <= 0 || n > kMaxInt32) return false;
      if (mem && mem->as_int64_ok) {
        if (mem->as_int64 < 0 || mem->as_int64 > kMaxInt32) return false;
        memnum = mem->as_int64;
        if (v.memory_factor > 1) { memnum *= v.memory_factor; if (memnum > kMaxInt32) return false; }
      }
      corenum = 100;
      if (core && core->as_int64_ok) {
        if (core->as_int64 < 0 || core->as_int64 > kMaxInt32) return false;
        corenum = (int32_t)core->as_int64;
      }
      mempnum = memnum == 0 ? 100 : 0;
      break;
    }
    case HS_REQ_MLU: {
      if (mem) {
        if (!mem->as_int64_ok || mem->as_int64 < 0 || mem->as_int64 > kMaxInt32 / factor) return false;
        memnum = mem->as_int64 * factor;
      }
      corenum = 100;
      if (core && core->as_int64_ok) corenum = (int32_t)core->as_int64;
      mempnum = memnum == 0 ? 100 : 0;
      break;
    }
    case HS_REQ_ASCEND: {
      if (n <= 0 || n > kMaxInt32) return false;
      if (mem) {
        if (mem->sign < 0) return false;                      // -1m has no int64 form: reject by sign first
        if (mem->as_int64_ok) {
          if (mem->as_int64 > kMaxInt32) return false;
          memnum = mem->as_int64;
          if (v.memory_factor > 1) { memnum *= v.memory_factor; if (memnum > kMaxInt32) return false; }
        }
      }
      corenum = 0;
      if (core) {
        if (core->value < 0 || core->value > kMaxInt32) return false;   // Quantity.Value(), rounded up
        corenum = (int32_t)core->value;
      }
      mempnum = memnum == 0 ? 100 : 0;
      break;
    }
    case HS_REQ_ILUVATAR: {                                   // iluvatar/device.go:201-254 (MemoryFactor 256); no range check on n
      if (mem) {
        if (!mem->as_int64_ok || mem->as_int64 < 0 || mem->as_int64 > kMaxInt32 / factor) return false;
        memnum = mem->as_int64 * factor;
      }
      corenum = 0;
      if (core && core->as_int64_ok) corenum = (int32_t)core->as_int64;
      mempnum = memnum == 0 ? 100 : 0;
      break;
    }
    case HS_REQ_COUNT:                                        // biren/device.go:141-165, vastai/device.go:143-167
      memnum = 0; mempnum = 100; corenum = 0;
      break;
    case HS_REQ_WHOLE_CARD:                                   // enflame/gcu.go:91-111
      memnum = 100; mempnum = 100; corenum = 100;
      break;
    case HS_REQ_METAX:                                        // metax/device.go:134-154
      memnum = 0; mempnum = 100; corenum = 100;
      break;
    case HS_REQ_AMD: {                                      // amd/device.go:217-266: Limits only, Requests are ignored
      cnt = lookup_limits(c, v.count_name);
      if (!cnt || !cnt->as_int64_ok) return false;
      if (cnt->as_int64 <= 0 || cnt->as_int64 > kMaxInt32) return false;
      r.nums = (int32_t)cnt->as_int64;
      mem = lookup_limits(c, v.memory_name);
      if (mem) {
        if (!mem->as_int64_ok || mem->as_int64 < 0 || mem->as_int64 > kMaxInt32) return false;
        memnum = mem->as_int64;
      }
      corenum = 100;                                          // an omitted core limit = every CU of each card
      core = lookup_limits(c, v.cores_name);
      if (core) {
        if (!core->as_int64_ok || core->as_int64 < 1 || core->as_int64 > 100) return false;
        corenum = (int32_t)core->as_int64;
      }
      mempnum = 0;
      r.key = v.common_word; r.type = v.common_word;
      r.memreq = (int32_t)memnum; r.mempct = mempnum; r.coresreq = corenum;
      return true;
    }
    case HS_REQ_MTHREADS: {                                   // mthreads/device.go:202-264 (MemoryFactor 512): per-card shares
      if (n <= 0 || n > kMaxInt32) return false;
      if (mem) {
        if (!mem->as_int64_ok || mem->as_int64 < 0 || mem->as_int64 > kMaxInt32 / factor) return false;
        memnum = mem->as_int64 * factor;
      }
      corenum = 0;
      if (core && core->as_int64_ok) corenum = (int32_t)core->as_int64;
      mempnum = memnum == 0 ? 100 : 0;
      memnum = (int32_t)memnum / (int32_t)n;                  // Memreq: int32(memnum) / int32(n)
      corenum = corenum / (int32_t)n;
      break;
    }
    case HS_REQ_METAX_SGPU:
    case HS_REQ_ENFLAME:                                      // handled before the count lookup (their own request shapes)
      return false;
    case HS_REQ_KUNLUN_VXPU: {                                // kunlun/vdevice.go:173-213: memory rounded up to 24576 / 49152 / 98304
      if (mem && mem->as_int64_ok) memnum = mem->as_int64 <= 24576 ? 24576 : (mem->as_int64 <= 49152 ? 49152 : 98304);   // trimMemory, :62-70
      corenum = (int32_t)(memnum * 100 / 98304);
      mempnum = 0;
      if (memnum == 0) { mempnum = 100; corenum = 100; memnum = 98304; }
      break;
    }
    default:
      return false;
  }
  r.key = v.common_word;
  r.type = v.common_word;
  r.nums = (int32_t)n;
  r.memreq = (int32_t)memnum;
  r.mempct = mempnum;
  r.coresreq = corenum;
  return true;
}

}  // namespace

extern "C" int32_t hs_generate_requests(const hs_vendor_resources* vendors, uint32_t n_vendors, const hs_container_resources* ctr,
                                        hs_container_request* out, uint32_t cap) {
  if ((!vendors && n_vendors) || !ctr || (!out && cap)) return HS_ERR_INVALID;
  uint32_t k = 0;
  for (uint32_t i = 0; i < n_vendors; i++) {
    hs_container_request r;
    if (!generate(vendors[i], ctr, r) || r.nums <= 0) continue;   // devices.go:660-663: only requests with Nums > 0 are kept
    if (k < cap) out[k] = r;
    k++;
  }
  return (int32_t)k;
}
