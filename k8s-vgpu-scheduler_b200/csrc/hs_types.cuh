// hs_types.cuh — plain structs shared by the host side (hs_engine.cu) and the
// sm_100a kernels (hs_kernels.cu).  Everything the device sees is integers:
// strings (types, UUIDs, policies, annotations) are resolved on the host.
#pragma once
#include <stdint.h>

#define HS_D 8            // device slots per node
#define HS_TILE 32        // nodes per inventory tile (one warp wide)
#define HS_MAX_E 16       // (container, vendor) request entries per pod
#define HS_MAX_C 16       // containers per pod
#define HS_MAX_ALLOC 64   // allocation rows one decision can produce
#define HS_SCORE_SLOTS 4  // distinct Ascend910* type keys per pod that contribute ScoreNode
#define HS_WIN 16         // ops decided per window: one inventory sweep and one grid barrier serve all of them (hs_window.cuh)
#define HS_WSLOTS 32      // nodes the window's resolver can track (one quarter warp each: the decide kernel's CTA is 256 threads)
#define HS_WMAX_TILES 4096 // inventories up to 131,072 nodes per GPU are decided in windows

// ---- inventory in HBM -------------------------------------------------------
// Structure-of-arrays in tiles of 32 nodes: for tile t, device slot d, lane n
//   dyn [ (t*HS_D + d)*HS_TILE + n ]  = {used, usedmem, usedcores, dflags}
//   stat[ (t*HS_D + d)*HS_TILE + n ]  = {count, totalmem, totalcore, packed}
// so a warp that owns one tile reads each (d) row as one coalesced 512 B line
// group with 128-bit loads.  32 B per device, 256 B per node.
//
// dflags: bit0 health (as getNodesUsage would hand it to Fit) | bit1 cordoned | bit2 present | bits3-5 mode id |
//         bit6 health as registered | bits8-15 NetworkID (255 = none) | bits16-23 MIG slices covered by the reservations
//         in use | bits24-31 number of cached pods that make the device fail closed (scheduler.go:788-841)
// packed: bits0-7 numa | bits8-13 type id | bits14-21 DeviceInfo.Index | bits22-26 vendor id | bits27-30 MIG geometry (0 = none)
#define DF_HEALTH 1u
#define DF_CORDON 2u
#define DF_PRESENT 4u
#define DF_MODE(x) (((x) >> 3) & 7u)
#define DF_NET(x) (((x) >> 8) & 255u)
#define DF_HEALTH_REG 64u
#define DF_SGPU 128u                // a metax sGPU device: bits 3-5 = its registered QoS policy (mode table id), bits 8-15 = its MetaLink zone,
                                    // bits 16-21 / 22-27 = online / offline pods on it, bit 28 + bits 29-31 = a QoS switch in flight and its target
#define SG_ON(x) (((x) >> 16) & 63u)
#define SG_OFF(x) (((x) >> 22) & 63u)
#define SG_JQ(x) (((x) >> 28) & 1u)
#define SG_JQID(x) (((x) >> 29) & 7u)
#define DF_MODE_MIG 2u              // index of "mig" in the handle's mode table (fixed: "", "hami-core", "mig", "mps")
#define DF_OCC(x) (((x) >> 16) & 255u)
#define DF_BAD(x) ((x) >> 24)
#define PK_GEOM(x) (((x) >> 27) & 15u)
#define PK_NUMA(x) ((int)((x) & 255u))
#define PK_TYPE(x) (((x) >> 8) & 63u)
#define PK_INDEX(x) ((int)(((x) >> 14) & 255u))
#define PK_VENDOR(x) (((x) >> 22) & 31u)

// node_flags: bits0-3 ndev | bit4 alive | bit5 pair scores uniform | bit6 hami-vnpu-core supported
#define NF_NDEV(x) ((int)((x) & 15u))
#define NF_ALIVE 16u
#define NF_PAIR_UNIFORM 32u
#define NF_VNPU_CORE 64u
#define NF_BASE_NV 128u     // some device of the node has Type == "NVIDIA" exactly: the key pads like a base type (score.go:71-79,246-250)
#define NF_BASE_MLU 256u    // … Type == "MLU"
#define NF_METAX_LOSS 512u  // the node has metax-tech.com/gpu.topology.losses: pair[0..31] = value for that many devices
#define NF_ALL_KUNLUN 2048u // every device of the node has a Type containing "kunlun" (kunlun ScoreNode returns 0 otherwise, device.go:157-160)
#define NF_METAX_SCORE 1024u // … gpu.topology.scores: pair[32..63]  (metax/device.go:185-214)

// vendor predicate ids (device-side switch, one code path, no dispatch layer)
enum { HV_NVIDIA = 0, HV_DCU = 1, HV_MLU = 2, HV_ASCEND0 = 3, HV_NONE = 31 };   // 5-bit ids: up to 31 registered vendors

// GPU-policy comparator (policy/gpu_policy.go:77-167), resolved on the host
enum { SORT_BINPACK = 0, SORT_SPREAD = 1, SORT_MUTEX = 2, SORT_NUMA_BINPACK = 3, SORT_NUMA_SPREAD = 4, SORT_CHAIN = 5 };
enum { CK_NONE = 0, CK_BINPACK = 1, CK_SPREAD = 2, CK_NUMA = 3 };

// entry flags
#define EF_ASCEND910 1u     // CommonWord has prefix "Ascend910": NetworkID topology + ScoreNode
#define EF_ASCEND910C 2u    // k.Type == "Ascend910C": module pairing
#define EF_MIGMODE_OK 4u    // pod's vgpu-mode annotation is absent or contains "mig" (isMIGRequest)
#define EF_NEURON 256u     // AWS Neuron's Fit (awsneuron/device.go:443-541): core bit masks, graphSelect for nums > 1, awsusageinfo state
#define EF_METAX 128u     // metax whole GPUs: EF_MTHREADS' gates (no uuid lists reach the device) + ScoreNode from the node's tables
#define EF_WHOLE_CARD 64u  // enflame GCU's Fit (enflame/gcu.go:144-183): health, type, "in use at all" = conflict; no uuid lists, no slot count
#define EF_MTHREADS 32u    // mthreads' Fit (mthreads/device.go:289-405): template gates + customFilterRule (one card per pod)
#define EF_AMD 16u         // amd's Fit (amd/device.go:279-371): whole-card memory by default, cores as a share of the card
#define EF_VASTAI 512u     // vastai: EF_COUNT_ONLY gates; on a node of dies (NetworkID = AIC) card exclusivity under mutex, pick by card, ScoreNode
#define EF_KUNLUN 1024u    // kunlun whole XPUs (kunlun/device.go:184-240, topo.go): graghSelect over list positions, slot-only usage, ScoreNode by wings
#define EF_VXPU 2048u      // with EF_KUNLUN: the shared "XPU" type (kunlun/vdevice.go:231-306): FitVXPU, full usage, no ScoreNode
#define EF_ENFLAME 4096u   // enflame shared vGCU (enflame/device.go:382-478): one DRS profile per request from the node's table, then the gates
#define EF_SGPU 8192u      // metax sGPU (metax/sdevice.go:332-495): filter by class / QoS / template gates, then prioritise; policy-neutral ScoreNode
#define EF_COUNT_ONLY 8u   // biren's Fit (biren/device.go:190-248): gates end after mutex; the pick has Usedmem 0, Usedcores = Coresreq

struct DevEntry {           // one ContainerDeviceRequest (devices.go:222-228), 64 B
  int32_t nums, memreq, mempct, coresreq;
  uint64_t type_contains;   // bit t: lower(type t) contains lower(k.Type)          (score.go:43-52)
  uint64_t type_equal;      // bit t: k.Type == type t exactly                     (gpu_policy.go:203)
  uint64_t type_allow;      // bit t: checkType verdict for type t                 (devices.go:719-738)
  uint32_t mode_allow;      // bit m: nvidia.com/vgpu-mode contains mode m         (nvidia/device.go:495-498)
  uint16_t vendor;          // HV_*; HV_NONE → "Device type not found"
  uint16_t flags;
  uint32_t use_off, nouse_off;   // sorted (node<<8 | mask8) words in the uuid pool: use-uuid / no-use-uuid lists
  uint16_t use_cnt, nouse_cnt;
  int8_t score_slot;        // ScoreNode accumulator (Ascend910*), -1 none
  uint8_t has_use;          // a non-empty use-uuid annotation exists
  uint8_t key_id;           // id of k.Type among the pod's distinct type keys (PodDevices key)
  uint8_t is_init;          // entry belongs to an init container
};

// pod flags
#define PF_NODE_SPREAD 1u   // node policy == "spread": pick the MIN score (node_policy.go:48-54)
#define PF_MUTEX 2u         // PolicyContains(gpuPolicy, "mutex")
#define PF_TOPOLOGY 4u      // PolicyContains(gpuPolicy, "topology-aware")  (NVIDIA)
#define PF_NUMA_ASSERT 8u   // nvidia.com/numa-bind parses true (assertNuma)
#define PF_ASCEND_HAMICORE 16u // huawei.com/vnpu-mode == "hami-core"
#define PF_NODE_BINPACK 64u // node policy == "binpack" (metax ScoreNode reads the policy string, metax/device.go:192-211)
#define PF_REL_MIG 128u   // OP_RELEASE of a known pod that touches MIG state: the release pool rows carry it in .w (bit0: one
                          // fail-closed count less, bit1: set the covered slices to bits 8-15)
#define PF_DRYRUN 32u     // decide and report the devices, leave the inventory untouched (filterSimulation, scheduler.go:1156-1199)

enum { OP_FILTER = 0, OP_RELEASE = 1, OP_NOP = 2 };

struct DevPod {             // one op of the stream, 96 B + entries
  uint32_t op;              // OP_*
  uint32_t flags;           // PF_*
  uint8_t sort_mode;        // SORT_*
  uint8_t chain[3];         // CK_* keys in order (SORT_CHAIN)
  uint8_t n_ctr, n_init, n_entries;
  uint8_t fast;             // exactly one request entry of the common shape: register-resident path (hs_fast.cuh)
  float w_slot, w_core, w_mem;   // float32(int64 weight)
  uint32_t entry_off;       // first DevEntry in the entry pool
  uint8_t ctr_begin[HS_MAX_C + 1]; // container c owns entries [ctr_begin[c], ctr_begin[c+1])
  uint8_t win_len;          // ops from this one on that may be decided as ONE window (hs_window.cuh); < 2 = decide it alone
  uint8_t sg_class;         // metax sGPU: bits 0-1 metax-tech.com/sgpu-app-class (0 absent, 1 online, 2 offline, 3 another value), bit 2 topology awareness on
  uint8_t sg_qos;           // … metax-tech.com/sgpu-qos-policy as a mode table id ("" = 0, a value the table does not hold = 7)
  // OP_RELEASE: subtract a committed allocation.
  int32_t rel_node;         // node slot, or -1 = take node and deltas from decision `rel_ref` of this batch
  int32_t rel_ref;
  int32_t rel_off;          // first of HS_D {slots, mem, cores, 0} int4 rows in the release pool
  int32_t q_idx[2];         // namespace quota rows (KernelArgs.quota) the NVIDIA [0] / MLU [1] Fit consults, -1 = none (quota.go:62-95)
  uint32_t alloc_base;      // FILTER: first DevAlloc row of this op (prefix sum of the ops' worst-case row counts: no cursor atomic)
  uint32_t pad4[5];
  DevEntry e0;             // the first (usually only) entry rides inside the op record: one 160 B prefetch per op
};
// OP_RELEASE of a pod the host knows (rel_node >= 0): the 8 per-device deltas ride in the op record, in the bytes
// of e0, so the owner needs no second fetch (the rows in rel_pool stay for the streaming and quota paths)
struct RelDelta { int32_t mem; int16_t cores; uint16_t slots; };
static_assert(sizeof(RelDelta) * HS_D == 64, "RelDelta rows overlay DevPod::e0");
static_assert(sizeof(DevEntry) == 64, "DevEntry layout");
static_assert(sizeof(DevPod) == 160, "DevPod layout: 96 B header + 64 B first entry, 16 B chunks for cp.async");

// numa-bind (nvidia/device.go:770-777) restarts Fit's collection whenever the NUMA id changes along the sorted list.
// With numa-bind the comparator is numa-major for every policy but "mutex" (policy/gpu_policy.go:101-112,139-147), so
// each NUMA id is ONE run and the verdict does not depend on the order inside it.  Under "mutex" (sorted by Used,
// descending, then NUMA id) the mutex gate only lets idle devices through; those sit together at the tail of the list,
// grouped by NUMA id, so each NUMA group of idle devices is one run as well and the devices in use before them can only
// reset a run, never add to it: the verdict is the same per-NUMA count
// (tests/test_oracle_manual_golden.py::test_numa_bind_verdict_under_mutex_is_a_per_numa_count_too).  Only a topology-aware
// request under the mutex comparator would need the real order — and the exact policy string "mutex" is never topology-aware.
#if defined(__CUDACC__)
#define HS_HD __host__ __device__
#else
#define HS_HD
#endif
HS_HD inline bool hs_numa_major(const DevPod& p) {
  return p.sort_mode == SORT_NUMA_BINPACK || p.sort_mode == SORT_NUMA_SPREAD || (p.sort_mode == SORT_CHAIN && p.chain[0] == CK_NUMA);
}
// Ascend910* on the lane-per-device path: the node's rank includes ScoreNode, which depends on the picks
HS_HD inline bool hs_is_net_scored(const DevPod& p) { return (p.e0.flags & EF_ASCEND910) != 0; }
HS_HD inline bool hs_verdict_needs_order(const DevPod& p) {
  if (hs_is_net_scored(p)) return true;
  return (p.flags & PF_NUMA_ASSERT) && p.e0.vendor == HV_NVIDIA && !hs_numa_major(p) &&
         !(p.sort_mode == SORT_MUTEX && !(p.flags & PF_TOPOLOGY));
}

// Dynamic MIG (nvidia/device.go:602-664, nvidia/mig_topology.go): what selectMigCandidate needs of one GPU's
// MigProfiles, prepared on the host — the candidates in its stable (MemoryMB, SliceCount) order, each with the placements
// of the FIRST profile of that name (findMigProfile) in selectMigPlacement's order (descending start for 1g./3g.).
// place = (start << 4) | size, start + size <= 8, so a placement is an 8-bit slice mask.
struct MigCand { int32_t mem; int16_t core; uint8_t n_place; uint8_t prof; uint8_t place[8]; };
struct MigGeom { uint32_t n_cand; uint32_t pad[3]; MigCand cand[16]; };
static_assert(sizeof(MigCand) == 16 && sizeof(MigGeom) == 272, "MIG geometry table layout");
HS_HD inline uint32_t hs_mig_mask(uint8_t place) { return (((1u << (place & 15u)) - 1u) << (place >> 4)) & 255u; }
// selectMigCandidate: -1 = none, else candidate | placement index << 4 | slice mask << 8
HS_HD inline int hs_mig_select(const MigGeom& g, uint32_t occupied, int memory) {
  for (uint32_t c = 0; c < g.n_cand && c < 16u; c++) {
    if (g.cand[c].mem < memory) continue;
    for (int p = 0; p < (int)g.cand[c].n_place && p < 8; p++) {
      const uint32_t m = hs_mig_mask(g.cand[c].place[p]);
      if (!(m & occupied)) return (int)c | (p << 4) | (int)(m << 8);
    }
  }
  return -1;
}

// enflame's DRS profiles of a node ride in the same table (PK_GEOM of its devices): one MigCand per candidate in
// selectProfileByRequest's order, mem = the profile's MiB (-1: a profile the reference rejects as ModeNotFit), core = its
// core share, place[0..3] = slice count, place[4..7] = GB (both int32, little endian).
HS_HD inline int hs_enf_i32(const MigCand& c, int at) {
  return (int)((uint32_t)c.place[at] | ((uint32_t)c.place[at + 1] << 8) | ((uint32_t)c.place[at + 2] << 16) | ((uint32_t)c.place[at + 3] << 24));
}
// selectProfileByRequest (enflame/device.go:498-531): candidate index or -1
HS_HD inline int hs_enf_select(const MigGeom& g, int memreq, int mempct, int coresreq) {
  const int n = (int)(g.n_cand < 16u ? g.n_cand : 16u);
  if (n == 0) return -1;
  if (mempct == 0) {                                   // direct: the slice count itself
    for (int c = 0; c < n; c++) if (hs_enf_i32(g.cand[c], 0) == memreq) return c;
    return -1;
  }
  const long long maxMemGB = hs_enf_i32(g.cand[n - 1], 4);
  long long req = 0;                                   // normalizeMemoryRequestToGB, :645-657
  if (memreq > 0) req = (maxMemGB > 0 && (long long)memreq > maxMemGB * 2) ? ((long long)memreq + 1023) / 1024 : memreq;
  for (int c = 0; c < n; c++) {
    if (req > 0 && (long long)hs_enf_i32(g.cand[c], 4) < req) continue;
    if (coresreq > 0 && (int)g.cand[c].core < coresreq) continue;
    return c;
  }
  return -1;
}

struct DevAlloc {           // one ContainerDevice, 8 B
  uint8_t entry;            // bits 0-3: index into the pod's entries (container + type); bits 4-7: MIG candidate (AL_MIG only)
  uint8_t dev;              // bits 0-2: device slot on the node; bit 3: a pick on a GPU in MIG mode; bits 4-6: its placement
  int16_t usedcores;
  int32_t usedmem;
};

#define AL_ENTRY(a) ((int)((a).entry & 15u))
#define AL_DEV(a) ((int)((a).dev & 7u))
#define AL_MIG(a) (((a).dev & 8u) != 0)
#define AL_MIG_CAND(a) ((int)((a).entry >> 4))
#define AL_MIG_PLACE(a) ((int)(((a).dev >> 4) & 7u))

struct DevDecision {        // 16 B header; rows live in the alloc pool
  int32_t status;           // 0 fit, 1 no fit, 2 release/nop done, <0 error
  int32_t node;             // node slot
  float score;
  uint32_t alloc_off;       // first DevAlloc
  uint32_t n_alloc;
  uint32_t fail_slot;       // no fit: slot in the fail buffer holding the per-node reasons (0xFFFFFFFF = not recorded)
  uint32_t pad[2];
};

// per-node failure record written only when a decision finds no node (common.go:46-53)
enum { FK_NONE = 0, FK_HIST = 1, FK_NODE_INSUFFICIENT = 2, FK_TYPE_NOT_FOUND = 3, FK_NOT_CANDIDATE = 4,
       FK_MIG_ADD_FAILED = 5,
       FK_HIST_ZALLOC = 6 };      // FK_HIST whose AllocatedCardsInsufficientRequest entry is present with a count of 0 (Ascend910C:
                                  // ascend/device.go:585-588 writes len(combination) into the map even when it is 0)   // AddResourceUsage returned an error (score.go:125-128): c0 = device slot, c1 = node slot
struct DevFail {            // 16 B
  uint32_t kind_cards;      // bits0-3 kind | bits4-31 cards denominator
  uint32_t c0;              // reasons 0-7, 4 bits each (HS_R_* order)
  uint32_t c1;              // reasons 8-15
  uint32_t pad;             // reasons 16-23 (metax sGPU's CardUnhealthy / CardAppClassMismatch / CardQosPolicyMismatch); FK_HIST only
};

// one 16 B slot per op: the arg-best key and the grid barrier that publishes it share an L2 sector
struct OpSync {
  unsigned long long best;      // packed argbest key (make_key), 0 = no node fits
  uint32_t arrived;             // CTAs that have folded their key in
  uint32_t pad;
};

struct KernelArgs {
  // inventory
  int4* dyn;
  const int4* stat;
  const uint32_t* node_flags;
  int4* node_tot;               // per node {flags, float bits of ComputeDefaultScore on the current usage, 0, 0}: kept current by every commit/release
  const int16_t* pair;          // [node][64]
  int32_t no_resident;          // test knob (HS_NO_RESIDENT=1): take the streaming mode even when the inventory would stay in registers
  const MigGeom* mig_tab;       // MIG geometries by PK_GEOM (entry 0 is empty); NULL while no GPU in MIG mode is registered
  int32_t n_nodes;              // slots (padded to HS_TILE)
  // candidate positions: NULL = every alive node, position == slot
  const int32_t* cand_pos;      // per slot: position in NodeNames or -1
  // ops
  const DevPod* pods;
  const DevEntry* entries;
  const uint32_t* uuid_pool;
  const int4* rel_pool;
  int32_t n_ops;                // ops [op_begin, n_ops) run in this launch
  int32_t op_begin;
  // results
  DevDecision* decisions;
  DevAlloc* allocs;
  uint32_t alloc_cap;
  int4* commit_delta;           // [op][HS_D] collapsed {slots, mem, cores, 0} of a committed decision
  OpSync* sync;                 // [op] {argbest key, CTA arrivals}, zeroed by the host
  DevFail* fail;                // [fail_cap][n_nodes]
  uint32_t fail_cap;
  uint32_t* counters;           // [0] alloc cursor, [1] fail cursor, [2] ops completed, [3] barrier arrive, [4] barrier epoch
  // multi-GPU: global position of this shard's slot 0 (argbest key carries global positions)
  int32_t pos_base;
  // fused multi-GPU arg-best (peer memory over NVLink): every rank's leader CTA stores its shard's key into each
  // peer's mailbox and polls its own; see xgpu_argbest in hs_kernels.cu.  world <= 1: unused.
  int32_t world, rank;
  unsigned long long* mbox_peer[8];   // mbox_peer[p] = rank p's mailbox, mapped into this process: [4 rows][8 ranks]
  unsigned long long* xresult;        // local: [4] epoch-tagged global winner, polled by every CTA of this GPU
  uint32_t epoch_base;                // barriers executed by earlier launches of this handle (epochs never restart)
  uint32_t* xerror;                   // set to 1 if a peer never arrived (bounded spin)
  // namespace quotas with a limit set: [row][4] = {memory limit x factor, memory used, core limit, core used}; a limit that
  // is not set is INT64_MAX.  Ops of such namespaces end with a second grid barrier (sync2) so the update is visible.
  long long* quota;
  OpSync* sync2;                // [op]
  unsigned long long* timing;   // diagnostic: 8 phase-clock accumulators written by block 0 (NULL = off)
  // windows (hs_window.cuh): per 32-node tile the nodes' arg-best keys sorted best first, per op of the window the fit
  // bits in that order and the tile's best fitting key; NULL = windows off
  unsigned long long* w_skey;   // [tile][32]
  uint32_t* w_rmask;            // [tile][HS_WIN]
  unsigned long long* w_head;   // [HS_WIN][n_tiles]
  uint8_t* w_rank;              // [node] rank of the node inside its tile's sorted order
  // diagnostic (HS_TRACE=1): per op of a launch [trace_ops][trace_stride] %globaltimer ns: column c = CTA c's arrival at the
  // op's barrier, column stride-2 = CTA of the winner (as a number), column stride-1 = when CTA 0 saw the barrier complete
  unsigned long long* trace;
  int32_t trace_ops, trace_stride;
  // chained launches of one pipelined batch: {ops of the batch completed so far, fail-buffer cursor} as the
  // predecessor left them (chain_in) and where this launch leaves its own (chain_out; the two alternate, so a CTA that
  // starts late never reads what its own launch wrote).  A launch whose op_begin differs from chain_in[0] does nothing:
  // its predecessor stopped early, the host drains the fail buffer and submits the rest again.
  const uint32_t* chain_in;
  uint32_t* chain_out;
  uint32_t* w_stats;            // [0] windows, [1] ops decided in windows, [2] stops on no-fit, [3] stops on a full tracker, [4] winners loaded late
};
