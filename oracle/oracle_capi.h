/*
 * oracle_capi.h — TEST INFRASTRUCTURE ONLY.  C entry points of the CPU oracle
 * (oracle.h) so tests/ and bench.py can drive it with ctypes.  It reuses the
 * plain input/output structs of include/hami_sched.h so the very same buffers
 * can be handed to the oracle and to the CUDA engine.
 */
#ifndef HAMI_ORACLE_CAPI_H_
#define HAMI_ORACLE_CAPI_H_
#include <stddef.h>
#include <stdint.h>
#include "../include/hami_sched.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_handle oracle_handle;

/* device.DeviceUsage literal (devices.go:106-125) + DeviceListsScore.Score */
typedef struct oracle_dev {
  const char* id;
  uint32_t    index;
  int32_t     used, count, usedmem, totalmem, totalcore, usedcores;
  const char* mode;
  int32_t     numa;
  const char* type;
  uint8_t     health;
  int32_t     network_id; /* -1 absent */
  float       score;
  const hs_mig_profile*   mig_profiles;   /* DeviceUsage.MigProfiles */
  uint32_t                n_mig_profiles;
  const hs_mig_placement* mig_in_use;     /* placements of DeviceUsage.MigAllocationsInUse */
  uint32_t                n_mig_in_use;
  const hs_enflame_info*  enflame;        /* CustomInfo of an enflame DRS device */
  const hs_metax_sgpu_info* sgpu;         /* CustomInfo of a metax sGPU device */
  int32_t                 online_pods, offline_pods;   /* DeviceUsage.PodInfos by metax-tech.com/sgpu-app-class */
} oracle_dev;

oracle_handle* oracle_create(const hs_config* cfg);
void    oracle_destroy(oracle_handle* o);
/* faithful=1: rebuild usage from the whole pod cache on every Filter like scheduler.go:744-863; threads: node fan-out */
void    oracle_set_mode(oracle_handle* o, int faithful, int threads);
int32_t oracle_node_upsert(oracle_handle* o, const hs_node_info* node);
int32_t oracle_node_remove(oracle_handle* o, const char* name);
int32_t oracle_usage_apply(oracle_handle* o, const char* pod_uid, const char* node, const char* uuid,
                           int32_t slots, int32_t usedmem, int32_t usedcores);
int32_t oracle_pod_delete(oracle_handle* o, const char* pod_uid);
/* calcscore (kunlun/topo.go:74-95) after addidx's dedup */
/* MetaxSDevices.ScoreNode (metax/sdevice.go:295-304) before OverrideScore's +-10000 weight */
float   oracle_sgpu_score(oracle_handle* o, const hs_alloc_entry* picks, uint32_t n_picks, const int32_t* zones,
                          const oracle_dev* prev, uint32_t n_prev, const hs_kv* annos, uint32_t n_annos);
int32_t oracle_snapshot_node_sgpu(oracle_handle* o, const char* node, int32_t* online, int32_t* offline, const char** pending, uint32_t cap);
float   oracle_kunlun_calc_score(const int32_t* prev, uint32_t n_prev, const int32_t* cur, uint32_t n_cur);
/* dynamic MIG: the informer side and the state getNodesUsage derives (see hs_pod_set_mig / hs_snapshot_node_mig) */
int32_t oracle_pod_set_mig(oracle_handle* o, const char* pod_uid, const hs_mig_reservation* res, uint32_t n);
int32_t oracle_snapshot_node_mig(oracle_handle* o, const char* node, uint32_t* occupied, uint8_t* health, uint32_t cap);
int32_t oracle_mig_select_placement(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                                    const char* profile, hs_mig_placement* out);
int32_t oracle_mig_can_place(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                             const char* const* requested, uint32_t n_req);
int32_t oracle_mig_select_candidate(const hs_mig_profile* profiles, uint32_t n, const hs_mig_placement* occupied, uint32_t n_occ,
                                    int32_t memory, int32_t* profile_index, hs_mig_placement* out);
/* podManager.GetPod (pods.go:146-155): node and collapsed per-device usage of one cached pod, summed per UUID;
 * returns the number of rows or HS_ERR_NOTFOUND.  Strings are valid until the next call. */
int32_t oracle_pod_get(oracle_handle* o, const char* pod_uid, const char** node_name, hs_pod_usage_row* rows, uint32_t cap);
int32_t oracle_filter(oracle_handle* o, const hs_pod* pod, const char* const* node_names, uint32_t n_names,
                      hs_filter_result* out);
int32_t oracle_snapshot_node(oracle_handle* o, const char* node, hs_device_usage_row* rows, uint32_t cap);
const char* oracle_last_error(const oracle_handle* o);
/* Run a whole op stream sequentially (Filter / DeletePod) and fold the results into the same FNV-1a-64
 * digest hs_batch_summary computes (see include/hami_sched.h).  seconds = wall time of the loop. */
int32_t oracle_run_ops(oracle_handle* o, const hs_op* ops, uint32_t n_ops, uint32_t* n_fit, uint32_t* n_nofit,
                       uint64_t* digest, double* seconds);

/* scoreNode (score.go:294-346) on an explicit NodeUsage, as Test_calcScore drives it.
 * list_policy NULL / numa_bind <0: derive from the pod like buildNodeUsage. status HS_FIT or HS_NOFIT (failed[0].reason). */
int32_t oracle_score_node(oracle_handle* o, const char* node_id, const oracle_dev* devs, uint32_t n,
                          const char* list_policy, int32_t numa_bind, const hs_pod* pod,
                          const hs_node_info* node_info, hs_filter_result* out);
/* fitInDevices (score.go:81-137) for ONE container's request map. Returns 1 fit / 0 no fit (reason in out->failed[0]). */
int32_t oracle_fit_in_devices(oracle_handle* o, const oracle_dev* devs, uint32_t n, const char* list_policy,
                              int32_t numa_bind, const hs_container* ctr, const hs_pod* pod,
                              const hs_node_info* node_info, hs_filter_result* out);
/* <vendor>.Fit on an explicit, already ordered device list. Returns 1/0; entries = tmpDevs[k.Type]. */
int32_t oracle_fit(oracle_handle* o, const char* vendor_key, const oracle_dev* devs, uint32_t n,
                   const hs_container_request* req, const hs_pod* pod, const hs_node_info* node_info,
                   hs_filter_result* out);

float   oracle_compute_score(const oracle_dev* d, const hs_container_request* reqs, uint32_t n,
                             int64_t w_slot, int64_t w_core, int64_t w_mem);
int32_t oracle_less(const char* policy, int32_t numa_bind, const oracle_dev* a, const oracle_dev* b);
void    oracle_sort(const char* policy, int32_t numa_bind, const oracle_dev* devs, uint32_t n, int32_t* order_out);
float   oracle_node_default_score(const oracle_dev* devs, uint32_t n);
int32_t oracle_check_uuid(const hs_kv* annos, uint32_t n, const char* id, const char* use_key, const char* nouse_key);
int32_t oracle_check_type(const hs_kv* annos, uint32_t n, const char* card_type, const char* use_key, const char* nouse_key);
int32_t oracle_parse_weights(const char* value, int64_t out3[3]);
/* reasons: HS_R_COUNT counts indexed by the HS_R_* enum */
int32_t oracle_gen_reason(const int32_t* counts, int32_t cards, char* buf, size_t cap);
/* CollapseInitContainerUsage (initContainer.go:36-114): entries in, collapsed entries out (idx field = Slots). */
int32_t oracle_collapse(const hs_alloc_entry* in, uint32_t n_in, uint32_t n_init, hs_alloc_entry* out, uint32_t cap);

/* QuotaManager (pkg/device/quota.go), keyed by namespace and (vendor word, which: 0 memory / 1 cores): the Go shim
 * maps "limits.<resource name>" to the vendor through GetResourceNames.  set = addQuotaLocked for one key,
 * del = delQuotaLocked; get returns 0 when the entry exists. */
int32_t oracle_quota_set(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t limit);
int32_t oracle_quota_del(oracle_handle* o, const char* ns, const char* vendor, int32_t which);
int32_t oracle_quota_get(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t* limit, int64_t* used, int32_t* limit_set);
/* test hooks for the vectors of pkg/device/quota_test.go: poke Used, and FitQuota with an explicit memory factor */
int32_t oracle_quota_add_usage(oracle_handle* o, const char* ns, const char* vendor, int32_t which, int64_t delta);
/* <vendor>.ScoreNode on one container of n_picks devices (metax/device.go:185-214 reads the node's two annotations) */
float oracle_vendor_score_node(oracle_handle* o, const char* vendor, uint32_t n_picks, const char* policy, const char* metax_losses,
                               const char* metax_scores);
int32_t oracle_fit_quota_scalar(oracle_handle* o, const char* ns, const char* vendor, int64_t memreq, int32_t factor, int64_t coresreq);
/* fitQuota (nvidia/device.go:666-692): tmp / allocated as flattened entries (type_key = PodDevices key). 1 = fits. */
int32_t oracle_fit_quota(oracle_handle* o, const char* ns, uint32_t n_init, const hs_alloc_entry* tmp, uint32_t n_tmp,
                         const hs_alloc_entry* allocated, uint32_t n_allocated, int32_t has_allocated, const char* vendor,
                         const char* dev_uuid, int64_t memreq, int64_t coresreq);

#ifdef __cplusplus
}
#endif
#endif
