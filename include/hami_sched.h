/*
 * hami_sched.h — C ABI of the B200-native HAMi Filter/Score engine.
 *
 * This is the drop-in boundary for ONE path of the reference scheduler
 * (soitun/k8s-vgpu-scheduler == Project-HAMi/HAMi v2.10.0): the block
 *   getNodesUsage + calcScore + sort.Sort/take-last + commit
 * inside Scheduler.Filter (pkg/scheduler/scheduler.go:1067-1154).  A Go
 * maintainer binds these entry points with cgo; every
 * argument is a caller-owned plain buffer that is copied before the call
 * returns, nothing calls back into the host language and no host pointer is
 * retained (cgo pointer rules).
 *
 * All paths in comments are relative to the reference tree root.
 *
 * Threading: a handle is thread-compatible, not thread-safe.  The reference
 * runs one scheduling cycle at a time; the cgo shim serialises calls with one
 * mutex, which also gives the deterministic order parity needs.
 *
 * Result memory returned through hs_filter_result / hs_batch_get is owned by
 * the handle and stays valid until the next call on that handle.
 */
#ifndef HAMI_SCHED_H_
#define HAMI_SCHED_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HS_MAX_DEVICES_PER_NODE 8   /* device slots per node held in HBM (reference asserts <=12 for stable sort, F6) */
#define HS_MAX_TYPES 64             /* distinct device model strings interned per handle */
#define HS_MAX_REQS_PER_POD 16      /* (container, vendor) request entries evaluated on the device per pod */
#define HS_MAX_CONTAINERS 16        /* init + app containers per pod */

/* ---- status codes ------------------------------------------------------ */
enum {
  HS_OK = 0,
  HS_FIT = 0,            /* a node was chosen and committed               */
  HS_NOFIT = 1,          /* no node fits: FailedNodes is filled, no commit (scheduler.go:1109-1114) */
  HS_PASSTHROUGH = 2,    /* pod requests no HAMi resource: every candidate passes (scheduler.go:1080-1087) */
  HS_FIT_REMOTE = 3,     /* sharded mode: a node on ANOTHER rank's shard was chosen; node_index = global position */
  HS_ERR_INVALID = -1,   /* malformed argument / unsupported shape        */
  HS_ERR_CUDA = -2,      /* CUDA runtime failure; see hs_last_error       */
  HS_ERR_UNSUPPORTED = -3,/* feature outside the built scope */
  HS_ERR_WEIGHTS = -4,   /* invalid hami.io/device-scoring-weights: whole Filter errors (score.go:361-364) */
  HS_ERR_NOTFOUND = -5
};

/* Fixed reason vocabulary of pkg/device/common/common.go:25-44, in this order. */
enum {
  HS_R_CardTypeMismatch = 0,
  HS_R_CardUuidMismatch,
  HS_R_CardTimeSlicingExhausted,
  HS_R_CardComputeUnitsExhausted,
  HS_R_CardInsufficientMemory,
  HS_R_CardInsufficientCore,
  HS_R_CardNotHealth,
  HS_R_CardCordoned,
  HS_R_NumaNotFit,
  HS_R_ExclusiveDeviceAllocateConflict,
  HS_R_CardNotFoundCustomFilterRule,
  HS_R_CardMigTopologyInfeasible,
  HS_R_AllocatedCardsInsufficientRequest,
  HS_R_ResourceQuotaNotFit,
  HS_R_ModeNotFit,
  HS_R_CardUnhealthy,             /* metax sGPU's own three (metax/sdevice.go:50-54) */
  HS_R_CardAppClassMismatch,
  HS_R_CardQosPolicyMismatch,
  HS_R_COUNT
};

/* ---- inventory --------------------------------------------------------- */

/* Dynamic MIG: device.MigPlacement / device.MigProfile (pkg/device/devices.go:50-53,79-90) as the
 * device plugin reports them for a GPU in MIG mode.  A placement with size 0 is refused (the reference's Fit
 * would visit such a GPU forever, nvidia/device.go:846-848); start + size <= 8, at most 16 profiles per GPU,
 * 8 placements per profile and 15 distinct geometries (profile lists) per handle.                              */
typedef struct hs_mig_placement {
  uint32_t start, size;
} hs_mig_placement;
typedef struct hs_mig_profile {
  const char*             name;        /* "1g.10gb", …                                  */
  int32_t                 memory_mb;
  int32_t                 core;
  uint32_t                slice_count;
  const hs_mig_placement* placements;
  uint32_t                n_placements;
} hs_mig_profile;

typedef struct hs_kv {
  const char* key;
  const char* value;
} hs_kv;

/* CustomInfo of one enflame shared-vGCU ("DRS") device as GetNodeDevices builds it (enflame/device.go:235-252):
 * minor, index, the node's profile map (profile name such as "3g.20gb" -> profile id) and maxSlice.               */
typedef struct hs_enflame_info {
  const char*  minor;
  const char*  index;
  int32_t      max_slice;
  const hs_kv* profiles;
  uint32_t     n_profiles;
} hs_enflame_info;

/* CustomInfo of a metax sGPU device (metax/protocol.go:96-116): the QoS policy the device runs ("", "best-effort",
 * "fixed-share", "burst-share") and its MetaLink zone (0 = none).                                                    */
typedef struct hs_metax_sgpu_info {
  const char* qos_policy;
  int32_t     link_zone;
} hs_metax_sgpu_info;

/* device.DeviceInfo (pkg/device/devices.go:127-144) as the register path hands
 * it to nodeManager.addNode (pkg/scheduler/nodes.go:76).                    */
typedef struct hs_device_info {
  const char* id;        /* UUID                                          */
  uint32_t    index;     /* DeviceInfo.Index                              */
  int32_t     count;     /* split count (slots)                           */
  int32_t     devmem;    /* MiB                                           */
  int32_t     devcore;   /* percent (or aiCore for Ascend)                */
  const char* type;      /* model string, e.g. "NVIDIA A100-SXM4-80GB"    */
  int32_t     numa;
  const char* mode;      /* "hami-core" | "mig" | "mps" | ""              */
  uint8_t     health;
  const char* vendor;    /* DevicesMap key == CommonWord(): "NVIDIA", "DCU", "MLU", "Ascend910B3", … */
  int32_t     network_id;/* CustomInfo["NetworkID"] (Ascend), -1 = absent */
  const hs_mig_profile* mig_profiles;   /* DeviceInfo.MIGProfiles (devices.go:141); read for mode "mig" only */
  uint32_t    n_mig_profiles;
  const hs_enflame_info* enflame;       /* HS_FIT_ENFLAME_VGCU devices only; NULL = no CustomInfo                */
  const hs_metax_sgpu_info* sgpu;       /* HS_FIT_METAX_SGPU devices only; NULL = no CustomInfo                  */
} hs_device_info;

/* device.NodeInfo (pkg/device/devices.go:170-174) plus the node annotations the
 * hot path reads.  Devices are listed vendor by vendor in sorted vendor-key
 * order, registration order inside a vendor (canonical rule 5). */
typedef struct hs_node_info {
  const char*           name;
  const hs_device_info* devices;
  uint32_t              n_devices;
  /* NVIDIA pair scores from annotation hami.io/node-nvidia-score
   * (nvidia/device.go:43,319-340): row-major n_devices x n_devices,
   * pair_scores[i*n+j] = DevicePairScore{ID=i}.Scores[j]; NULL = none (0). */
  const int32_t*        pair_scores;
  const char*           device_cordon;  /* annotation hami.io/device-cordon (nvidia/device.go:59), NULL = absent */
  int8_t                hami_vnpu_core; /* annotation hami-vnpu-core: -1 absent, 1 == "true", 0 any other value (ascend/device.go:463-467) */
  /* annotations metax-tech.com/gpu.topology.losses / .scores (metax/device.go:38-39,185-214): JSON {"<devices>": value};
   * NULL = absent.  Read by the HS_FIT_METAX vendor's ScoreNode; such a node must hold that vendor's devices only. */
  const char*           metax_losses;
  const char*           metax_scores;
} hs_node_info;

/* ---- pod --------------------------------------------------------------- */

/* device.ContainerDeviceRequest (pkg/device/devices.go:222-228), i.e. the
 * output of <vendor>.GenerateResourceRequests; `key` is the map key it is
 * stored under in ContainerDeviceRequests (devices.go:659,677).             */
typedef struct hs_container_request {
  const char* key;
  const char* type;
  int32_t     nums;
  int32_t     memreq;
  int32_t     mempct;    /* MemPercentagereq; 101 = unset (nvidia/device.go:551) */
  int32_t     coresreq;
} hs_container_request;

typedef struct hs_container {
  const hs_container_request* reqs;
  uint32_t                    n_reqs;
  const char*                 name;   /* Container.Name; read by the one PatchAnnotations that keys on it (enflame's
                                       * assigned-containers, enflame/device.go:268-297); NULL = "container-<index>"  */
} hs_container;

/* The parts of *corev1.Pod that Filter reads: UID, namespace, annotations,
 * and Resourcereqs(pod) (devices.go:639-691): init containers first.       */
typedef struct hs_pod {
  const char*         uid;
  const char*         ns;
  const char*         name;
  const hs_kv*        annotations;
  uint32_t            n_annotations;
  const hs_container* containers;
  uint32_t            n_containers;
  uint32_t            n_init;     /* len(pod.Spec.InitContainers) */
} hs_pod;

/* ---- results ----------------------------------------------------------- */

/* device.ContainerDevice (pkg/device/devices.go:205-216), flattened with the
 * container index and the PodDevices map key it belongs to.  Entries are raw
 * (un-collapsed), ordered by type key, then container, then pick order.    */
typedef struct hs_alloc_entry {
  uint32_t    container;
  const char* type_key;  /* PodDevices key == request k.Type */
  const char* uuid;
  int32_t     idx;
  int32_t     usedmem;
  int32_t     usedcores;
  /* CustomInfo["migProfile"] / ["migPlacement"] of a pick on a GPU in MIG mode (nvidia/device.go:653-657): the
   * reservation AddResourceUsage made; usedmem / usedcores are then the profile's.  NULL = not a MIG pick.       */
  const char* mig_profile;
  uint32_t    mig_start, mig_size;
} hs_alloc_entry;

/* Per failed node: reason string exactly as common.GenReason renders it
 * (common/common.go:46-53) or the node-level reason ("NodeInsufficientDevice",
 * "node unregistered", …).                                                 */
typedef struct hs_failed_node {
  const char* node;
  const char* reason;
} hs_failed_node;

/* What Scheduler.Filter produces (scheduler.go:1117-1153): the chosen
 * policy.NodeScore{NodeID, Devices, Score} (policy/node_policy.go:27-33), the
 * annotation map it patches onto the pod, or FailedNodes.                  */
typedef struct hs_filter_result {
  int32_t               status;      /* HS_FIT | HS_NOFIT | HS_PASSTHROUGH | <0 */
  int32_t               node_index;  /* position in the candidate list      */
  const char*           node_name;
  float                 score;
  const hs_alloc_entry* entries;
  uint32_t              n_entries;
  const hs_kv*          annotations; /* without hami.io/vgpu-time / predicate-time values (set to "") */
  uint32_t              n_annotations;
  const hs_failed_node* failed;
  uint32_t              n_failed;
} hs_filter_result;

/* One row of the usage cache: DeviceUsage{Used, Usedmem, Usedcores}.        */
typedef struct hs_device_usage_row {
  int32_t used;
  int32_t usedmem;
  int32_t usedcores;
} hs_device_usage_row;

/* ---- configuration ----------------------------------------------------- */

/* One Ascend VNPUConfig (pkg/device/ascend/device.go VNPUConfig; defaults in
 * pkg/scheduler/config/config.go:340-460).  Templates are only used to name the
 * vNPU template in the huawei.com/<word> annotation (ascend/device.go:69-79,221-253). */
typedef struct hs_ascend_template {
  const char* name;
  int64_t     memory;
} hs_ascend_template;

typedef struct hs_ascend_config {
  const char*               common_word;        /* "Ascend910B3", "Ascend910C", "Ascend310P", … */
  int64_t                   memory_allocatable;
  int64_t                   memory_capacity;
  const hs_ascend_template* templates;          /* sorted by memory ascending at load, like InitDevices */
  uint32_t                  n_templates;
} hs_ascend_config;

/* cmd/scheduler/main.go:70-71 flags + the vendor registry of
 * pkg/scheduler/config/config.go:107-251.  NVIDIA, DCU and MLU are always
 * registered; Ascend words as listed.                                       */
/* A vendor beyond NVIDIA / DCU / MLU / Ascend, registered next to the built-in ones; fit_kind names
 * whose Fit it runs (seven kinds below).
 *   HS_FIT_TEMPLATE    iluvatar (pkg/device/iluvatar/device.go:267-371): the gate list of hygon's Fit — health, type,
 *                      uuid, slots, mutex, memory, cores, exclusive, compute units; checkType = key equals the word
 *                      (:190-195); one registration per configured CommonWord ("BI-V100", "BI-V150", "MR-V100", …)
 *   HS_FIT_COUNT_ONLY  biren (pkg/device/biren/device.go:190-248): health, type, uuid, slots, mutex only; the pick
 *                      records Usedmem 0 and Usedcores = Coresreq as requested (no clamp)
 * PatchAnnotations of iluvatar, biren and amd writes EncodePodSingleDevice under the two keys (iluvatar :143-154, biren :167-177,
 * amd :120-131).
 *   HS_FIT_AMD        amd (pkg/device/amd/device.go:279-371): health, type (use/no-use type lists, :165-186), uuid,
 *                      slots, mutex; Memreq <= 0 means the whole card's memory; Coresreq > 0 is a share of the card's
 *                      cores (Totalcore * Coresreq / 100, clamped to [1, Totalcore]), a whole-memory request without
 *                      cores takes all cores; the pick records the DEVICE's type string, not k.Type (:341-346)      */
enum { HS_FIT_TEMPLATE = 0, HS_FIT_COUNT_ONLY = 1, HS_FIT_AMD = 2,
       /* mthreads (pkg/device/mthreads/device.go:289-405): the template gates plus customFilterRule (:266-276) — every
        * device an earlier container of the pod received must be the candidate itself, reason
        * "CardNotFoundCustomFilterRule"; PatchAnnotations (:155-177) writes only the allocated key plus
        * mthreads.com/gpu-index, mthreads.com/predicate-node and mthreads.com/predicate-time (returned empty)         */
       HS_FIT_MTHREADS = 3,
       /* enflame GCU, whole cards (pkg/device/enflame/gcu.go:144-183): health, type, and a card in use at all is an
        * "ExclusiveDeviceAllocateConflict"; no uuid lists, no slot count; the pick records Memreq and Coresreq as asked */
       HS_FIT_WHOLE_CARD = 4,
       /* metax whole GPUs, "Metax-GPU" (pkg/device/metax/device.go:223-334): the template gates without uuid lists
        * (checkUUID is constant true) plus mthreads' customFilterRule; plain PatchAnnotations (:91-103); ScoreNode
        * (:185-214) adds 2000 - losses[n] under node policy "binpack" and 2000 - scores[n] under "spread", n = the
        * devices the pod received, from hs_node_info.metax_losses / metax_scores                                    */
       HS_FIT_METAX = 5,
       /* AWS Neuron (pkg/device/awsneuron/device.go:443-541).  Devcore / usedcores are core BIT MASKS.  One device:
        * health, type, uuid, slots, mutex, then popcount(Totalcore) - popcount(Usedcores) >= Coresreq; the pick's
        * Usedcores is the mask addCoreUsage (:367-389) hands out, tracked per pod in CustomInfo["awsusageinfo"].
        * Several devices: graphSelect (:406-441) — a free, healthy run of list positions (any start on an "inf" ring
        * node, aligned groups of 4 / 8 / 16 elsewhere), each position mapped to the device with that Index; the picks
        * take the whole card (Totalmem, Totalcore); failure reads "NumaNotFit".  hs_device_info.network_id carries
        * CustomInfo["AWSNodeType"]: 1 = the instance type contains "inf" or "Inf", 0 = any other, -1 = absent.
        * Annotations: the allocated key only; AWS_NEURON_IDS, NEURON_RESOURCE_TYPE, NEURON_ALLOC_TIME, NEURON_ALLOCATED and
        * aws.amazon.com/predicate-node depend on the container's resource limits (:200-243), which hs_pod does not
        * carry — the caller's own PatchAnnotations adds them from hs_filter_result.entries.                          */
       HS_FIT_NEURON = 6,
       /* vastai (pkg/device/vastai/device.go:228-327).  Card mode: biren's gate list.  DIE mode (the node's devices say
        * CustomInfo["DeviceStrategy"] = "die"; hs_device_info.network_id carries CustomInfo["AIC"], the card a die sits
        * on, as a small integer that keeps the AIC strings' order): under mutex an idle die on a card with a die in use
        * is a conflict too (:241-250,283-297); Fit collects every fitting die and picks by card — all of them when
        * exactly `nums`, the first one when nums == 1, else cards by die count (descending; equal counts in ascending
        * AIC, canonical rule 7), dies inside a card in walk order (:306-320,336-372); ScoreNode = largest share of one
        * card among a container's dies, unweighted (:181-216).  A node must be all-die or all-card.                  */
       HS_FIT_VASTAI = 7,
       /* kunlun whole XPUs (pkg/device/kunlun/device.go:184-240, kunlun/topo.go).  Fit = graghSelect over the node's
        * kunlun devices in LIST order: positions 0-3 are the left wing, 4-7 the right; a device fits when it is idle and
        * healthy and passes BOTH uuid annotation pairs (hami.io/use-xpu-uuid / no-use-xpu-uuid and
        * baidu.com/use-gpuuuid / nouse-gpuuuid).  8 cards: all of them; 1, 2 or 4: the wing whose free count is the
        * tightest match (1 → 1,3,2,4; 2 → 2,3,4; 4 → 4), else an inter-group pair (2) or interconnect group (4) by card
        * Index; anything else does not fit.  Failure: the number of cards a uuid list turned away, else "1/n NumaNotFit".
        * The picks take the whole card (Totalmem, Totalcore) but AddResourceUsage only counts the slot (:179-182).
        * ScoreNode (:154-177): 0 unless every device of the node is a kunlun card, else 3000 / 2000 / 1000 / 0 by how
        * the "bubble" count of half-used wings changes (calcscore, topo.go:74-95).  Annotations: the allocated key and
        * BAIDU_COM_DEVICE_IDX = the card indices of the last container with devices (:93-110).  Card Index < 32.       */
       HS_FIT_KUNLUN = 8,
       /* kunlun shared XPUs, type "XPU" (pkg/device/kunlun/vdevice.go:231-306): HS_FIT_KUNLUN's graghSelect with FitVXPU —
        * healthy, Memreq + Usedmem <= Totalmem, and a card already in use only takes requests whose Memreq equals the
        * card's average (Usedmem / Used); under the mutex policy only idle cards.  No slot count.  The pick records the
        * request's Memreq / Coresreq, usage is added in full, ScoreNode is 0, both annotation keys are written.  Failure:
        * under mutex the number of in-use cards that would otherwise fit ("ExclusiveDeviceAllocateConflict"), else the
        * cards a uuid list turned away, else "1/n NumaNotFit".                                                          */
       HS_FIT_KUNLUN_VXPU = 9,
       /* enflame shared vGCU, type "Enflame" (pkg/device/enflame/device.go:382-478).  Every request is first turned into
        * ONE profile of the node (selectProfileByRequest, :498-531): the candidates are the profiles the node's devices
        * advertise (hs_enflame_info; "<slices>g.<GB>gb"), sorted by (slices, GB) — equal ones in ascending name — with a core
        * share of ceil(slices * 100 / maxSlice); MemPercentagereq 0 ("direct", a enflame.com/drs-gcu request) takes the
        * profile whose slice count equals Memreq, anything else the first profile that covers the memory (values above
        * twice the largest profile are read as MiB and rounded up to GB) and the core share asked.  No profile:
        * "1/n ModeNotFit".  Then health, type, uuid (enflame.com/use-gpuuuid / nouse-gpuuuid), slots, mutex, the
        * profile's memory (GB * 1024) and core share (when Totalcore > 0); the pick records those and the profile;
        * AddResourceUsage adds the slice count to Used.  hs_alloc_entry.mig_profile carries the profile name and
        * mig_start its slice count.  Annotations (:254-300): the allocated key, enflame.com/gcu-assigned = "false",
        * gcu-assigned-time (returned empty), gcu-assigned-index / -id / -minor and gcu-request-size of the first
        * container with a device, assigned-containers = {"<container name>": {"allocated":false,"request":<slices>,
        * "profileID":…,"profileName":…}} (hs_container.name).                                                          */
       HS_FIT_ENFLAME_VGCU = 10,
       /* metax sGPU, type "Metax-SGPU" (pkg/device/metax/sdevice.go:332-495).  Filter, then prioritise.  A device passes
        * when it is healthy ("CardUnhealthy"), its application class agrees with the pod's
        * metax-tech.com/sgpu-app-class (a device in use runs either online/offline pods or plain ones, and an offline pod
        * never joins a device with online pods: "CardAppClassMismatch"), — unless the pod is online — its QoS policy
        * agrees with metax-tech.com/sgpu-qos-policy (an idle device and an exclusive request always pass; the device's
        * policy is the one a previous Filter is switching it to, if any, else the registered one:
        * "CardQosPolicyMismatch"), then uuid lists (metax-tech.com/use-gpuuuid / nouse-gpuuuid), slots, mutex, memory
        * and — unless online — cores, exclusivity and exhaustion as in the template.  Fewer passing devices than asked:
        * no fit.  More: an online pod takes the devices with the fewest online, then offline, pods (devices that only run
        * offline pods last); a request for whole cards (Coresreq 100, no app class) takes a MetaLink zone that holds
        * enough of them (the smallest such zone; equal sizes in ascending zone), else fills up across zones, then with
        * unlinked devices; everything else the first ones in list order.  An online pod's picks record 0 cores.
        * ScoreNode (:295-304) is policy neutral — OverrideScore multiplies it by 10000, or -10000 under node policy
        * "spread" (policy/node_policy.go:82-88): online pods lose 100 per offline pod on a picked device that runs only
        * offline pods, else 20 per online and 2 per offline pod; whole-card pods with topology awareness (the vendor's
        * `option`, overridden by metax-tech.com/sgpu-topology-aware) score 10 x the MetaLink pairs among their cards minus
        * the pairs the node loses (:669-...).  Annotations: both hami keys, metax-tech.com/gpu-devices-allocated (JSON),
        * metax-tech.com/predicate-time (returned empty).  The engine keeps the class counts and the pending QoS switches
        * per device; hs_node_upsert of a device whose registered policy equals its pending one retires the switch
        * (JitteryQosCache.Sync, metax/qos.go:37-59).  Pods reach the cache with their class through hs_filter or
        * hs_pod_ingest (the annotations are read there); hs_usage_apply adds classless pods.                             */
       HS_FIT_METAX_SGPU = 11 };
typedef struct hs_vendor_config {
  const char* common_word;       /* DevicesMap key == k.Type == DeviceInfo.DeviceVendor                     */
  int32_t     fit_kind;          /* HS_FIT_*                                                                  */
  const char* use_uuid_key;      /* CheckUUID annotations, e.g. "birentech.com/use-gpuuuid"; NULL = none      */
  const char* nouse_uuid_key;
  const char* to_allocate_key;   /* InRequestDevices[word]; NULL = "hami.io/<word>-devices-to-allocate"       */
  const char* allocated_key;     /* SupportDevices[word];   NULL = "hami.io/<word>-devices-allocated"         */
  const char* use_type_key;      /* HS_FIT_AMD: "amd.com/use-gputype" / "amd.com/nouse-gputype"; NULL = no lists */
  const char* nouse_type_key;
  int32_t     option;            /* HS_FIT_METAX_SGPU: 1 = MetaxConfig.TopologyAware (sgpuTopologyAware); else 0          */
} hs_vendor_config;

typedef struct hs_config {
  const char*             node_scheduler_policy;  /* NULL = "binpack" */
  const char*             gpu_scheduler_policy;   /* NULL = "spread"  */
  const hs_ascend_config* ascend;
  uint32_t                n_ascend;
  uint8_t                 ascend_hami_vnpu_core;  /* VNPUs.HamiVnpuCore global default */
  int32_t                 cuda_device;            /* ordinal; -1 = current device      */
  uint32_t                reserve_nodes;          /* initial HBM capacity in nodes (grows on demand) */
  const hs_vendor_config* vendors;                /* further vendors with a plain Fit; at most 31 vendors in total */
  uint32_t                n_vendors;
} hs_config;

typedef struct hs_handle hs_handle;

/* Trace operations for hs_run_batch: the pod stream a scheduler sees. */
enum { HS_OP_FILTER = 0, HS_OP_DELETE = 1 };
typedef struct hs_op {
  int32_t       kind;
  const hs_pod* pod;   /* FILTER: the pod; DELETE: only uid is read */
} hs_op;

/* ---- entry points ------------------------------------------------------ */

/* Replaces newScheduler + config.InitDevicesWithConfig for this path.
 * Fails (returns NULL, message via hs_last_error(NULL)) when no CUDA device
 * is usable: there is no CPU fallback.                                      */
hs_handle* hs_create(const hs_config* cfg);
void       hs_destroy(hs_handle* h);
const char* hs_last_error(const hs_handle* h);

/* nodeManager.addNode / rmNodeDevices (pkg/scheduler/nodes.go:76-148): insert
 * or replace a node's static inventory.  Usage already committed on devices
 * whose UUID survives the update is kept.                                   */
int32_t hs_node_upsert(hs_handle* h, const hs_node_info* node);
int32_t hs_node_remove(hs_handle* h, const char* name);
int32_t hs_node_count(const hs_handle* h);

/* podManager.AddPod fed by the informer (scheduler.go:137-171, pods.go:57-97):
 * add an already-collapsed allocation (restart replay, pods bound elsewhere).
 * `slots` 0 counts as 1 (scheduler.go:782).                                 */
int32_t hs_usage_apply(hs_handle* h, const char* pod_uid, const char* node,
                       const char* uuid, int32_t slots, int32_t usedmem, int32_t usedcores);

/* podManager.DelPod + usage release (scheduler.go:226-253, pods.go:157-167). */
int32_t hs_pod_delete(hs_handle* h, const char* pod_uid);

/* Scheduler.Filter (scheduler.go:1067): one pod, sweep + select + commit.
 * node_names == NULL means "every registered node, in registration order";
 * otherwise the ExtenderArgs.NodeNames candidate list (unknown names fail with
 * "node unregistered", scheduler.go:846-853).                               */
int32_t hs_filter(hs_handle* h, const hs_pod* pod,
                  const char* const* node_names, uint32_t n_names,
                  hs_filter_result* out);

/* The same decisions for a whole pod stream, committed one at a time ON THE
 * DEVICE in stream order (a persistent kernel, one grid barrier per decision),
 * so assignments equal n sequential hs_filter/hs_pod_delete calls.
 * Candidates are every registered node.  Returns the number of ops run.     */
int32_t hs_run_batch(hs_handle* h, const hs_op* ops, uint32_t n_ops);
/* The same, without waiting for the end of the batch: a long batch runs as a pipeline inside the library (request
 * compile of chunk c+1 | device run of chunk c | usage-cache update of chunk c-1) and the call returns as soon as
 * that pipeline is started, so the caller can read the results of the first chunks (hs_batch_get, hs_batch_summary;
 * both wait for the op they read) while the device still decides the later ones.  `ops` and everything it points to
 * must stay valid until the batch has finished: hs_batch_wait(h, HS_BATCH_ALL), a complete hs_batch_summary, or any
 * other call on the handle (each of them first lets the pipeline finish).  No callback enters the caller's code and no
 * caller pointer is kept beyond the batch (cgo rules).  Short batches simply run to completion inside the call.     */
int32_t hs_run_batch_async(hs_handle* h, const hs_op* ops, uint32_t n_ops);
#define HS_BATCH_ALL 0xffffffffu
/* Blocks until result i of the running batch can be read; HS_BATCH_ALL: until the whole batch has finished. */
int32_t hs_batch_wait(hs_handle* h, uint32_t i);
/* Result i of the last hs_run_batch (DELETE ops yield status HS_OK, no node). */
int32_t hs_batch_get(hs_handle* h, uint32_t i, hs_filter_result* out);

/* Walk every result of the last hs_run_batch exactly as a caller would (node, score, the
 * annotation map of scheduler.go:1124-1130, or FailedNodes) and fold it into one FNV-1a-64
 * digest; also counts fits / no-fits.  The consumer a throughput measurement needs, and the
 * "checksum of checksums" parity property at full size.  Layout of the digest: per op in order,
 * status (4 B LE); fit: node name, score bits (4 B LE), then key\0value\0 per annotation in
 * key order; no fit: node\0reason\0 per candidate in candidate order.  digest == NULL: every
 * result is still materialised (what a caller reading the batch does) but not hashed.          */
int32_t hs_batch_summary(hs_handle* h, uint32_t* n_fit, uint32_t* n_nofit, uint64_t* digest);

/* The reference undoes the commit when the annotation PATCH fails
 * (scheduler.go:1140-1147): podManager.DelPod + quota RmUsage.              */
int32_t hs_rollback(hs_handle* h, const char* pod_uid);

/* Usage cache dump for one node, rows in device-slot order (metrics / parity;
 * the reference's equivalent is InspectAllNodesUsage, scheduler.go:630-639). */
int32_t hs_snapshot_node(hs_handle* h, const char* node, hs_device_usage_row* rows, uint32_t cap);

/* ---- sharded mode: one process per GPU, contiguous node blocks, NCCL allreduce-max per decision ----
 * (no counterpart in the reference, which runs one scheduler replica).  Rank 0 creates the
 * 128-byte NCCL unique id and ships it to the others (e.g. torch.distributed broadcast); every rank then
 * calls hs_shard_init with the global candidate position of its first node.  Afterwards every rank must
 * call hs_run_batch with the SAME ops; a decision is committed by the rank that owns the winning node and
 * reported as HS_FIT_REMOTE by the others.                                                             */
int32_t hs_shard_unique_id(void* out128);
int32_t hs_shard_init(hs_handle* h, int32_t rank, int32_t world, const void* unique_id128, int64_t pos_base);
/* Fused path: the arg-best merge happens INSIDE the persistent decide kernel over NVLink peer memory (each
 * rank's leader CTA stores its key into every peer's mailbox and polls its own), so a decision costs no
 * kernel launch and no NCCL call.  hs_shard_mailbox returns this rank's 64-byte CUDA IPC handle; gather all
 * ranks' handles (rank order, 64 bytes each) and pass them to hs_shard_connect on every rank.  Up to 8 ranks
 * on one node.  Without it (or with HS_SHARD_NCCL=1) hs_run_batch uses the NCCL launch-per-decision path.  */
int32_t hs_shard_mailbox(hs_handle* h, void* out64);
int32_t hs_shard_connect(hs_handle* h, const void* handles64, int32_t world);

/* ---- request extraction: device.Resourcereqs for one container (pkg/device/devices.go:639-691)
 * over <vendor>.GenerateResourceRequests (nvidia/device.go:520-600, hygon/device.go:168-239,
 * cambricon/device.go:250-305, ascend/device.go:303-378).  A resource.Quantity arrives as the three views the
 * reference reads: AsInt64() (+ ok), Value() (rounded up) and Sign().  Pure host code, needs no handle.            */
typedef struct hs_quantity {
  const char* name;         /* resource name, e.g. "nvidia.com/gpumem"      */
  int64_t     as_int64;     /* Quantity.AsInt64() value                     */
  uint8_t     as_int64_ok;  /* … and its ok flag                            */
  int64_t     value;        /* Quantity.Value()                             */
  int8_t      sign;         /* Quantity.Sign()                              */
  uint8_t     has_suffix;   /* Quantity.String() contains a letter (a unit): read by HS_REQ_METAX_SGPU only */
} hs_quantity;
typedef struct hs_container_resources {
  const hs_quantity* limits;   uint32_t n_limits;     /* ctr.Resources.Limits   */
  const hs_quantity* requests; uint32_t n_requests;   /* ctr.Resources.Requests */
} hs_container_resources;
enum { HS_REQ_NVIDIA = 0, HS_REQ_DCU = 1, HS_REQ_MLU = 2, HS_REQ_ASCEND = 3,
       HS_REQ_ILUVATAR = 4,   /* iluvatar/device.go:201-254: memory_factor 256, cores default 0                  */
       HS_REQ_COUNT = 5,      /* biren/device.go:141-165, vastai/device.go:143-167, kunlun/device.go:129-152: the count only */
       HS_REQ_AMD = 6,        /* amd/device.go:217-266: Limits only; cores = percent of the card, default 100     */
       HS_REQ_MTHREADS = 7,   /* mthreads/device.go:202-264: memory_factor 512; memory and cores divided by count */
       HS_REQ_WHOLE_CARD = 8, /* enflame/gcu.go:91-111: Memreq 100, MemPercentagereq 100, Coresreq 100            */
       HS_REQ_METAX = 9,      /* metax/device.go:134-154: Memreq 0, MemPercentagereq 100, Coresreq 100            */
       HS_REQ_KUNLUN_VXPU = 11, /* kunlun/vdevice.go:173-213: count + memory trimmed up to 24576 / 49152 / 98304 MiB, cores =
                               * the memory's share of a card in percent; no memory = the whole card                 */
       HS_REQ_METAX_SGPU = 13,/* metax/sdevice.go:222-293: Limits only; cores default 100; memory in Gi unless the quantity carries a
                               * unit (hs_quantity.has_suffix), then MiB                                                  */
       HS_REQ_ENFLAME = 12,   /* enflame/device.go:322-366: count_name (enflame.com/drs-gcu) > 0 = a direct slice request (Nums 1,
                               * Memreq = the count, MemPercentagereq 0); else memory_name / cores_name by spec (Nums 1,
                               * Memreq, Coresreq, MemPercentagereq -1); Quantity.Value() of Limits, then Requests        */
       HS_REQ_NEURON = 10     /* awsneuron/device.go:287-334: by device (count_name) or by NeuronCore (cores_name, split
                               * over devices); default_cores = the vendor object's coresPerAWSNeuron                */
};
typedef struct hs_vendor_resources {
  const char* common_word;             /* becomes the request's key and Type                                  */
  int32_t     kind;                    /* HS_REQ_*: whose GenerateResourceRequests                             */
  const char* count_name;              /* ResourceCountName, e.g. "nvidia.com/gpu"                             */
  const char* memory_name;             /* ResourceMemoryName                                                   */
  const char* memory_percentage_name;  /* NVIDIA only                                                          */
  const char* cores_name;              /* ResourceCoreName; NULL / "" = none (Ascend)                          */
  int32_t     default_memory;          /* NVIDIA DefaultMemory                                                 */
  int32_t     default_cores;           /* NVIDIA DefaultCores                                                  */
  int32_t     memory_factor;           /* MemoryFactor (MLU: 256)                                              */
} hs_vendor_resources;
/* Fills out[] with the container's requests whose Nums > 0 (key == type == common_word, pointers borrowed from
 * `vendors`), in vendor order; returns how many there are (may exceed cap).                                   */
int32_t hs_generate_requests(const hs_vendor_resources* vendors, uint32_t n_vendors, const hs_container_resources* ctr,
                             hs_container_request* out, uint32_t cap);

/* ---- ingest: the data formats that keep the HBM inventory true ---------------------------
 * The reference learns a node's devices from register annotations and a pod's committed allocation from its
 * "<vendor>-devices-allocated" annotations; these entry points take the annotation STRINGS.                */
enum { HS_REG_CSV = 0,    /* device.DecodeNodeDevices  (pkg/device/devices.go:317-381): DCU, iluvatar      */
       HS_REG_JSON = 1,   /* device.UnMarshalNodeDevices (pkg/device/devices.go:445-449): NVIDIA, Ascend, …;
                           * for both annotation formats capacity_cores > 0 overwrites every device's Devcore, which is
                           * what biren/device.go:84 and vastai/device.go:86 do with 100 after the decode              */
       /* no annotation at all: the devices are derived from the node's CAPACITY (capacity_cores / capacity_memory =
        * AsInt64 of Status.Capacity[<cores resource>] / [<memory resource>]) and named after the node:               */
       HS_REG_CAPACITY_MLU = 2,      /* cambricon/device.go:209-232: one card per 100 cores, <node>-cambricon-mlu-<i>   */
       HS_REG_CAPACITY_MTHREADS = 3, /* mthreads/device.go:130-153: one card per 16 cores, <node>-mthreads-<i>          */
       HS_REG_CAPACITY_METAX = 4,    /* metax/device.go:67-89: capacity_cores = the GPU count, <node>-metax-<i>, Count 1;
                                      * `annotation` / `pair_scores` = the node's metax-tech.com/gpu.topology.losses /
                                      * .scores values or NULL (hs_node_register hands them to hs_node_info)            */
       HS_REG_CAPACITY_KUNLUN = 6,   /* kunlun/device.go:64-91: capacity_cores = the XPU count, <node>-kunlun-<i>, Count 100,
                                      * 98304 MiB, the upper half of the cards on NUMA 1                                */
       HS_REG_CAPACITY_GCU = 9,      /* enflame/gcu.go:67-89: capacity_cores = the node's enflame.com/gcu count, <node>-GCU-<i>,
                                      * whole cards: Count 1, Devmem 100, Devcore 100                                    */
       HS_REG_METAX_SGPU = 8,        /* metax/sdevice.go:506-526 + protocol.go:96-116: `annotation` = the node's
                                      * metax-tech.com/node-gpu-devices JSON ([]MetaxSDeviceInfo); Index = position             */
       HS_REG_ENFLAME_DRS = 7,       /* enflame/device.go:198-252: `annotation` = the node's enflame.com/gcu-drs-capacity JSON
                                      * {"devices":[{"index","minor","capacity"}],"profiles":{name:id}}; <node>-enflame-drs-<i>,
                                      * Count = capacity, Devmem = the largest profile's GB * 1024, Devcore 100            */
       HS_REG_CAPACITY_NEURON = 5    /* awsneuron/device.go:156-198: capacity_cores = the DEVICE count, capacity_memory = the
                                      * node's total core count, capacity_aux = the vendor object's coresPerAWSNeuron so far
                                      * (0 = unknown: total / count, which the caller then keeps), `annotation` = the node's
                                      * node.kubernetes.io/instance-type label or NULL; <node>-AWSNeuron-<i>, Count = cores
                                      * per device, Devcore = their bit mask, network_id as HS_FIT_NEURON documents it       */
};

typedef struct hs_vendor_registration {
  const char* vendor;       /* CommonWord(); becomes DeviceInfo.DeviceVendor of every decoded device          */
  int32_t     format;       /* HS_REG_*                                                                      */
  const char* annotation;   /* value of the vendor's node register annotation (e.g. hami.io/node-nvidia-register) */
  const char* pair_scores;  /* value of hami.io/node-nvidia-score, NULL = absent (nvidia/device.go:319-340)   */
  const char* node_name;    /* HS_REG_CAPACITY_*: the node's name (device IDs are built from it); hs_node_register fills it in */
  int64_t     capacity_cores;
  int64_t     capacity_memory;
  int64_t     capacity_aux; /* HS_REG_CAPACITY_NEURON only */
} hs_vendor_registration;

/* One ContainerDevice of a decoded / collapsed PodDevices (devices.go:205-216): uuid NULL marks an EMPTY
 * container at (key, container) — DecodePodDevices keeps those to preserve the container index mapping.    */
typedef struct hs_pod_device {
  const char* key;          /* PodDevices map key (the vendor word of the checklist)                         */
  uint32_t    container;
  const char* uuid;
  const char* type;
  int32_t     usedmem;
  int32_t     usedcores;
  int32_t     slots;        /* 0 in raw decoded entries; set by the collapse functions                       */
} hs_pod_device;

/* <vendor>.GetNodeDevices without the node object (nvidia/device.go:302-345, hygon/device.go:131-151,
 * ascend/device.go:197-215): decode one vendor's register annotation (+ pair scores).  Pure host code, needs no
 * handle and no GPU.  Outputs point into thread-local storage, valid until the next hs_decode_* call on this
 * thread.  *pair_scores is row-major n x n (NULL when reg->pair_scores is NULL).  Errors: HS_ERR_INVALID with the
 * reference's error text in hs_last_error(NULL); an empty device list is HS_ERR_NOTFOUND ("no gpu found on node"). */
int32_t hs_decode_node_devices(const hs_vendor_registration* reg, const hs_device_info** devices, uint32_t* n_devices,
                               const int32_t** pair_scores);
/* device.DecodePodDevices (devices.go:526-556): checklist = (PodDevices key, annotation name) pairs, e.g.
 * SupportDevices {"NVIDIA": "hami.io/vgpu-devices-allocated"}.  Entries sorted by key, then container.      */
int32_t hs_decode_pod_devices(const hs_kv* checklist, uint32_t n_checklist, const hs_kv* annotations, uint32_t n_annotations,
                              const hs_pod_device** devices, uint32_t* n_devices);
/* mode 0: device.CollapseInitContainerUsage (initContainer.go:36-114); 1: AppContainersOnlyDeviceUsage (:118-157).
 * Output per key: container 0, devices sorted by UUID, slots set.                                            */
int32_t hs_collapse_pod_devices(int32_t mode, uint32_t n_init, const hs_pod_device* raw, uint32_t n_raw,
                                const hs_pod_device** devices, uint32_t* n_devices);

/* One node of the register loop (scheduler.go:469-533, health handshake excluded): decode every vendor's
 * annotation, skip vendors that fail to decode or report no device, and upsert the node with the devices that
 * remain (vendors in sorted key order, registration order inside a vendor).  No device left = the node is
 * removed (rmNodeDevices).  Returns the number of devices registered, or a negative status.                 */
int32_t hs_node_register(hs_handle* h, const char* node, const hs_vendor_registration* regs, uint32_t n_regs,
                         const char* device_cordon, int8_t hami_vnpu_core);
/* onAddPod (scheduler.go:137-171) / restart replay: DecodePodDevices over the registered vendors'
 * "*-devices-allocated" annotations, CollapseInitContainerUsage, podManager.AddPod (pods.go:57-97: an existing
 * pod's devices are replaced unless its init-container share was already released).  `node` = the value of
 * hami.io/vgpu-node; it must be registered (HS_ERR_NOTFOUND otherwise: ingest nodes first).                  */
int32_t hs_pod_ingest(hs_handle* h, const char* pod_uid, const char* ns, const char* node, const hs_kv* annotations,
                      uint32_t n_annotations, uint32_t n_init);

/* Dynamic MIG, the informer side (scheduler.go:757-842): the decoded hami.io/vgpu-mig-allocations annotation of a
 * pod the cache holds (nvidia.DecodeMigAllocations, nvidia/mig_allocations.go:78-112), in annotation order.  Filter
 * itself caches the pod WITHOUT it (args.Pod is not patched in place, scheduler.go:1135-1139), so every GPU in MIG
 * mode it picked fails closed ("MIG Pod lacks a matching profile/placement reservation") until the informer delivers
 * the patched pod: hs_pod_ingest decodes the annotation itself, hs_pod_set_mig is the same step for callers that
 * feed the cache through hs_usage_apply.  n = 0 clears the pod's reservations.                                      */
typedef struct hs_mig_reservation {
  const char* gpu_uuid;
  const char* profile;
  uint32_t    start, size;
} hs_mig_reservation;
int32_t hs_pod_set_mig(hs_handle* h, const char* pod_uid, const hs_mig_reservation* res, uint32_t n);
/* nvidia.DecodeMigAllocations (nvidia/mig_allocations.go:78-112) on its own: pure host code, output in thread-local
 * storage until the next hs_decode_* call on this thread; HS_ERR_INVALID with the reason in hs_last_error(NULL) for
 * malformed JSON, an incomplete entry, a partial runtime identity or a duplicate (containerIndex, deviceIndex).    */
int32_t hs_decode_mig_allocations(const char* json, const hs_mig_reservation** out, uint32_t* n);
/* What getNodesUsage would hand Fit for one node's MIG state: per device the occupied-slice mask (bit s = slice s
 * is covered by a reservation in use) and whether the device is healthy after the fail-closed rules.  Rows in the
 * node's device order; returns the number of devices.                                                           */
int32_t hs_snapshot_node_mig(hs_handle* h, const char* node, uint32_t* occupied, uint8_t* health, uint32_t cap);
/* The metax sGPU bookkeeping of one node, per device in the node's device order: the online / offline pods on it
 * (DeviceUsage.PodInfos by metax-tech.com/sgpu-app-class) and the QoS policy a previous Filter is switching it to
 * (JitteryQosCache; NULL = none; the strings belong to the handle).  Returns the number of devices.                 */
int32_t hs_snapshot_node_sgpu(hs_handle* h, const char* node, int32_t* online, int32_t* offline, const char** pending, uint32_t cap);

/* ---- namespace quota: device.QuotaManager, pkg/device/quota.go ------------------------------
 * The NVIDIA and MLU Fit refuse a device when the pod's collapsed total would push the namespace past a configured
 * limit (fitQuota, nvidia/device.go:666-692,784-788; cambricon/device.go:351-376,439-443; reason
 * "ResourceQuotaNotFit").  hs_quota_set / hs_quota_del = AddQuota / DelQuota for ONE managed key of a ResourceQuota
 * (quota.go:243-289): the caller maps "limits.<resource name>" to the vendor whose GetResourceNames() lists it and to
 * which = 0 (memory) or 1 (cores).  Used is maintained by every commit, release, hs_pod_ingest / hs_pod_shrink /
 * hs_pod_delete (AddUsage / RmUsage / ReplaceUsage, quota.go:119-196; a removal clamps at zero) for every namespace,
 * limit or not.  hs_quota_memory_factor = ResourceNames.MemoryFactor (default 1, MLU 256).  Pods of a namespace with a
 * limit take the general evaluator and end with a second grid barrier; not available on the sharded path.          */
int32_t hs_quota_set(hs_handle* h, const char* ns, const char* vendor, int32_t which, int64_t limit);
int32_t hs_quota_del(hs_handle* h, const char* ns, const char* vendor, int32_t which);
int32_t hs_quota_get(hs_handle* h, const char* ns, const char* vendor, int32_t which, int64_t* limit, int64_t* used, int32_t* limit_set);
int32_t hs_quota_memory_factor(hs_handle* h, const char* vendor, int32_t factor);
/* onUpdatePod once every init container has succeeded (scheduler.go:200-222): usage shrinks to
 * AppContainersOnlyDeviceUsage, once (podManager.UpdatePodDevice).                                           */
int32_t hs_pod_shrink(hs_handle* h, const char* pod_uid, const hs_kv* annotations, uint32_t n_annotations, uint32_t n_init);

/* podManager.GetPod (pkg/device/pods.go:146-155): what the pod cache holds for one pod — the node it is bound to and,
 * per device of that node it occupies, the collapsed {slots, memory, cores} the next usage rebuild would add
 * (scheduler.go:782-785).  Returns the number of rows written (devices with non-zero usage), HS_ERR_NOTFOUND for a pod
 * the cache does not hold.  `node_name` and the uuid strings belong to the handle (valid until the node changes).
 * Used by metrics and by parity harnesses that rebuild the same cache in a second scheduler.                       */
typedef struct hs_pod_usage_row {
  const char* uuid;
  int32_t     slots, usedmem, usedcores;
} hs_pod_usage_row;
int32_t hs_pod_get(hs_handle* h, const char* pod_uid, const char** node_name, hs_pod_usage_row* rows, uint32_t cap);

/* ---- simulation: Scheduler.filterSimulation, pkg/scheduler/scheduler.go:1088-1090,1156-1199 -----
 * hs_filter_dry: Filter against the handle's inventory that releases nothing, commits nothing and remembers
 * nothing (same result record as hs_filter).
 * hs_filter_simulate: the cluster-autoscaler path — the candidate nodes arrive WITH the request as node objects;
 * each is rebuilt from its register annotations with zero usage (buildTransientNodeInfo, :692-711; a node that
 * yields no device fails with "node unregistered") in a transient inventory on the same GPU, then scored dry.
 * The result belongs to the handle and is valid until its next call.                                          */
typedef struct hs_sim_node {
  const char*                   name;
  const hs_vendor_registration* regs;
  uint32_t                      n_regs;
  const char*                   device_cordon;
  int8_t                        hami_vnpu_core;
} hs_sim_node;
int32_t hs_filter_dry(hs_handle* h, const hs_pod* pod, const char* const* node_names, uint32_t n_names, hs_filter_result* out);
int32_t hs_filter_simulate(hs_handle* h, const hs_pod* pod, const hs_sim_node* nodes, uint32_t n_nodes, hs_filter_result* out);

/* ---- the allocation-free entry point (it has no counterpart in the reference) --------------------
 * For callers that cannot call cudaMalloc when they need a scan: a plugin whose device memory is a fixed
 * pool carved up at start-up, or a latency-critical path that must not enter the driver's allocator.
 * Scans the handle's resident inventory for `pod` (one request entry, no UUID use/no-use list; no
 * commit) and writes 16 bytes to the caller's device buffer: the winning packed key (bits 63-32 ordered
 * score, bits 31-8 node position + 1, 0 = no node fits), then the number of CTAs used.  All scratch lives
 * in one static device array per GPU (about 8.5 KB, held until the process exits): at most 1024 CTAs and
 * 4,194,304 nodes, larger inputs are refused with HS_ERR_UNSUPPORTED.  The first call on a device resolves
 * the array's address — make it before any stream capture.  Calls share the array: run them one after
 * the other, never overlapping on the GPU.  The call is synchronous (it returns once the scan has finished
 * and the result is in the caller's buffer) and it uses the handle's request-compile state like hs_filter
 * does, so never run it concurrently with another call on the same handle.
 * `cuda_stream` NULL = the engine's own stream.                                                         */
int32_t hs_sweep_static(hs_handle* h, const hs_pod* pod, void* out_device16, void* cuda_stream);

/* ---- measurement / diagnostic hooks (no counterpart in the reference) ----------------------- */
/* The inventory scan in isolation: compile `pod`, run the sweep-only kernel `iters` times over the
 * whole inventory (no commit); average device ms per sweep, the winning candidate position, its score. */
int32_t hs_sweep_probe(hs_handle* h, const hs_pod* pod, uint32_t iters, double* ms_avg, int32_t* node_index, float* score);
/* When each Filter decision of the last hs_run_batch became final on the device (%globaltimer ns; 0 = not a Filter). */
/* test hook: the dynamic records of one node as the DEVICE holds them ({used, usedmem, usedcores, dflags} x 8) next to the
 * host mirror's — after a batch the two must be identical (the kernel's commits / releases and the host's replay of them). */
int32_t hs_debug_device_rows(hs_handle* h, const char* node, int32_t device_rows[32], int32_t mirror_rows[32]);
int32_t hs_debug_decision_ns(hs_handle* h, uint64_t* out, uint32_t cap);
/* Cumulative SM clocks block 0 spent per phase of the decide loop (needs HS_TIMING=1 at hs_create). */
int32_t hs_debug_timing(hs_handle* h, unsigned long long out[8]);
/* All 16 phase clocks (8..11: inside the windows) and the window counters: windows, ops decided in windows, stops on a
 * no-fit op, stops on a full tracker, winners loaded in the middle of a window. */
int32_t hs_debug_window_stats(hs_handle* h, unsigned long long timing16[16], uint32_t stats8[8]);
/* Cumulative host seconds: request compile, device run + copies, result absorb, summary. */
int32_t hs_debug_host_seconds(hs_handle* h, double out[4]);
/* HS_TRACE=1: when every CTA arrived at the barrier of each of the first 512 ops of the last launch (%globaltimer ns),
 * [ops][stride]: column c = CTA c, column stride-2 = winner's CTA + 1, column stride-1 = CTA 0 saw the barrier complete. */
int32_t hs_debug_trace(hs_handle* h, uint64_t* out, uint32_t cap, uint32_t* ops, uint32_t* stride);
/* Cumulative bytes the batch path copied host->device [0] and device->host [1], counted at the copies themselves. */
int32_t hs_debug_copy_bytes(hs_handle* h, uint64_t out[2]);
/* The host parser behind hs_node_info.metax_losses / metax_scores on its own (parseMetaxAnnos, metax/device.go:168-183):
 * the table's entry for `index` as hs_node_upsert would store it; 0 for a missing key and for any document error. */
int64_t hs_debug_metax_table(const char* json, int64_t index);
/* Self-check of the branch-free float32 divide against the IEEE divide on n operand pairs; mismatches. */
long long hs_debug_divcheck(unsigned long long n, unsigned long long seed);

/* Device time (ms, CUDA events on the engine's stream) and kernel launches of
 * the last hs_filter / hs_run_batch, for bench.py.                          */
double   hs_last_device_ms(const hs_handle* h);
uint64_t hs_launch_count(const hs_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* HAMI_SCHED_H_ */
