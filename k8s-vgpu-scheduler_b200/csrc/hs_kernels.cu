// hs_kernels.cu — sm_100a kernels of the HAMi Filter/Score engine.
//
// hs_decide_kernel is ONE persistent cooperative kernel that runs a whole stream of scheduling
// ops in order.  Per FILTER op (reference: Scheduler.Filter, pkg/scheduler/scheduler.go:1067-1154):
//   1. the op's 160 B request record is staged in shared memory (cp.async prefetch of the next op
//      overlaps the current one);
//   2. every node is evaluated by the quarter warp that owns it — one lane per device slot, the
//      inventory held in registers (hs_fast.cuh) — or, for the uncommon request shapes, by one
//      thread running the general evaluator (hs_eval.cuh).  When the cluster is larger than the
//      resident lanes, owners stream their nodes from the SoA inventory with coalesced 128-bit loads;
//   3. warp-shuffle + shared-memory arg-best, then ONE 16-byte L2 slot per op carries both the
//      atomicMax of the packed (score, position) key and the grid barrier's arrival count;
//   4. the owner of the winning node commits the collapsed usage (CollapseInitContainerUsage +
//      AddPod's effect on the next usage rebuild) and writes the decision record.  Ownership of a
//      node is static, so a commit is only ever read back by the lanes that wrote it: no second
//      barrier, and the next op's sweep starts immediately everywhere else.
// A decision therefore sees every earlier commit, exactly like the reference's sequential Filter
// calls, and costs one barrier instead of a kernel launch.
//
// Integer/compare work on CUDA cores: there is no dense contraction on this path, so no tensor
// cores.
#include <cuda_runtime.h>

#include "hs_eval.cuh"
#include "hs_fast.cuh"
#include "hs_window.cuh"

// Inventory layout (node-major structure-of-arrays): dyn[node*8 + slot] = {used, usedmem, usedcores,
// dflags}, stat[node*8 + slot] = {count, totalmem, totalcore, packed}.  The 8 lanes of a node read
// 128 contiguous bytes of each array, a warp (4 nodes) 512 B: fully coalesced 128-bit loads.
__device__ __forceinline__ void load_node(const KernelArgs& ka, int node, uint32_t nflags, NodeWork& w) {
  const int4* dy = ka.dyn + (size_t)node * HS_D;
  const int4* st = ka.stat + (size_t)node * HS_D;
  int4 a[HS_D], b[HS_D];
#pragma unroll
  for (int d = 0; d < HS_D; d++) { a[d] = dy[d]; b[d] = __ldg(st + d); }
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    w.used[d] = a[d].x; w.mem[d] = a[d].y; w.cores[d] = a[d].z; w.dflags[d] = (uint32_t)a[d].w;
    w.count[d] = b[d].x; w.totmem[d] = b[d].y; w.totcore[d] = b[d].z; w.packed[d] = (uint32_t)b[d].w;
  }
  w.ndev = NF_NDEV(nflags);
  w.order = 0x76543210u;
}

__device__ __forceinline__ void load_dev(const KernelArgs& ka, int node, int slot, DevRegs& n) {
  const int4 a = ka.dyn[(size_t)node * HS_D + slot];
  const int4 b = __ldg(ka.stat + (size_t)node * HS_D + slot);
  n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
  n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
}

// CollapseInitContainerUsage (pkg/device/initContainer.go:36-114) folded into the usage
// the next rebuild would see (scheduler.go:782-785): per device slot {slots, mem, cores}.
__device__ void collapse_rows(const DevEntry* ent, const DevAlloc* al, uint32_t n, int4* delta) {
  for (int d = 0; d < HS_D; d++) delta[d] = make_int4(0, 0, 0, 0);
  // groups are (type key, device); visit each group once at its first row
  for (uint32_t i = 0; i < n; i++) {
    const int key = ent[AL_ENTRY(al[i])].key_id, dev = AL_DEV(al[i]);
    bool first = true;
    for (uint32_t j = 0; j < i; j++)
      if (ent[AL_ENTRY(al[j])].key_id == key && AL_DEV(al[j]) == dev) { first = false; break; }
    if (!first) continue;
    // a pick on a GPU in MIG mode: the cached pod carries no reservation yet, so the GPU fails closed (.w bit 0)
    if (AL_MIG(al[i]) && ent[AL_ENTRY(al[i])].vendor == HV_NVIDIA) delta[dev].w = 1;      // (enflame picks use the same spare bits for their profile)
    int imem = 0, icores = 0, islots = 0, amem = 0, acores = 0, aslots = 0;
    for (uint32_t j = i; j < n; j++) {
      if (ent[AL_ENTRY(al[j])].key_id != key || AL_DEV(al[j]) != dev) continue;
      if (ent[AL_ENTRY(al[j])].is_init) {
        if (al[j].usedmem > imem) imem = al[j].usedmem;
        if ((int)al[j].usedcores > icores) icores = (int)al[j].usedcores;
        islots = 1;
      } else {
        amem += al[j].usedmem;
        acores += (int)al[j].usedcores;
        aslots++;
      }
    }
    int slots = islots > aslots ? islots : aslots;
    if (slots < 1) slots = 1;
    delta[dev].x += slots;
    delta[dev].y += imem > amem ? imem : amem;
    delta[dev].z += icores > acores ? icores : acores;
  }
}

// countPodDevices + addUsageLocked / rmUsageLocked (quota.go:97-117,136-150,169-184) for the rows the Fit path reads:
// the pod's collapsed {slots, mem, cores} per device, summed per vendor; a removal clamps Used at zero.
__device__ __forceinline__ void quota_apply(const KernelArgs& ka, const DevPod& pod, int node, const int4* delta, int sign) {
  long long mem[2] = {0, 0}, core[2] = {0, 0};
  for (int d = 0; d < HS_D; d++) {
    const uint32_t vendor = PK_VENDOR((uint32_t)ka.stat[(size_t)node * HS_D + d].w);
    const int qv = vendor == HV_NVIDIA ? 0 : (vendor == HV_MLU ? 1 : -1);
    if (qv >= 0) { mem[qv] += delta[d].y; core[qv] += delta[d].z; }
  }
  for (int v = 0; v < 2; v++) {
    if (pod.q_idx[v] < 0) continue;
    long long* row = ka.quota + (size_t)pod.q_idx[v] * 4;
    const long long m = __ldcg(row + 1) + sign * mem[v], c = __ldcg(row + 3) + sign * core[v];
    __stcg(row + 1, m < 0 ? 0 : m);
    __stcg(row + 3, c < 0 ? 0 : c);
  }
}

// dflags after a change of the MIG bookkeeping (delta.w of a commit / release row): bit 0 = one more (sign +1) or one
// less cached pod that makes the device fail closed, bit 1 = the covered slices become bits 8-15
__device__ __forceinline__ uint32_t mig_apply(uint32_t df, int w, int sign) {
  if ((w >> 2) & 3) {          // metax sGPU: one online (1) / offline (2) pod more or less on the device
    const int sh = ((w >> 2) & 3) == 1 ? 16 : 22;
    int c = (int)((df >> sh) & 63u) + sign;
    c = c < 0 ? 0 : (c > 63 ? 63 : c);
    return (df & ~(63u << sh)) | ((uint32_t)c << sh);
  }
  if (w & 2) df = (df & ~(255u << 16)) | (((uint32_t)w >> 8) & 255u) << 16;
  if (w & 1) {
    int bad = (int)DF_BAD(df) + sign;
    bad = bad < 0 ? 0 : (bad > 255 ? 255 : bad);
    df = (df & 0x00ffffffu) | ((uint32_t)bad << 24);
  }
  return (DF_BAD(df) == 0u && (df & DF_HEALTH_REG)) ? (df | DF_HEALTH) : (df & ~DF_HEALTH);
}

__device__ __forceinline__ void store_usage(const KernelArgs& ka, int node, const int4* delta, int sign) {
  int4* dy = ka.dyn + (size_t)node * HS_D;
  for (int d = 0; d < HS_D; d++) {
    if (delta[d].x == 0 && delta[d].y == 0 && delta[d].z == 0 && delta[d].w == 0) continue;
    int4 v = dy[d];
    v.x += sign * delta[d].x; v.y += sign * delta[d].y; v.z += sign * delta[d].z;
    if (delta[d].w) v.w = (int)mig_apply((uint32_t)v.w, delta[d].w, sign);
    dy[d] = v;
  }
}

// Recompute ComputeDefaultScore of one node from HBM and publish it in node_tot (one thread; used where the
// node is not register-resident).  The streamed sweeps read the score from there instead of re-deriving it.
__device__ void refresh_node_score(const KernelArgs& ka, int node) {
  NodeWork w;
  load_node(ka, node, ka.node_flags[node], w);
  w.ndev = HS_D;   // absent slots hold zeros
  ka.node_tot[node].y = __float_as_int(node_default_score(w));
}

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t > v ? t : v;
  }
  return v;
}

__device__ __forceinline__ void pack_fail(const EvalOut& out, DevFail& f) {
  f.kind_cards = (out.fkind & 15u) | (out.fcards << 4);
  uint32_t c0 = 0, c1 = 0;
  const bool hist = out.fkind == FK_HIST || out.fkind == FK_HIST_ZALLOC;
  if (hist) {
    for (int r = 0; r < 8; r++) c0 |= (out.reasons[r] & 15u) << (4 * r);
    for (int r = 8; r < 16; r++) c1 |= (out.reasons[r] & 15u) << (4 * (r - 8));
  }
  uint32_t c2 = 0;
  if (hist)
    for (int r = 16; r < 24; r++) c2 |= (out.reasons[r] & 15u) << (4 * (r - 16));
  f.c0 = c0; f.c1 = c1; f.pad = c2;
}
// FK_MIG_ADD_FAILED names a device of the node in its text: the node slot rides along (fcards holds the device slot)
__device__ __forceinline__ void pack_fail_node(const EvalOut& out, DevFail& f, int node) {
  pack_fail(out, f);
  if (out.fkind == FK_MIG_ADD_FAILED) f.c1 = (uint32_t)node;
}

// Arg-best + grid barrier in one 16-byte L2 slot per op: {best key, arrivals}.  Warp 0 of every CTA
// max-reduces the CTA's warp keys with shuffles; its lane 0 folds the result in with atomicMax, then
// release-adds the CTA's arrival and polls the arrival count with acquire loads (slot_poll): every
// CTA's max is ordered before its arrival by the release, the acquire that reads the full count orders
// the read of the key after all of them.  The launch is cooperative, so every CTA is resident.  Returns the winning key to the CTA.
// (An all-to-all ring of per-CTA slots polled by a whole warp was measured slower: 157 pollers x 5
// lines per round hot-spot the L2 slices and delay the arrivals themselves.)
// Multi-GPU extension (one process per GPU, nodes sharded in contiguous blocks): the CTAs of a GPU first
// combine through the local slot exactly as above; CTA 0's lane 0 (the rank's leader) then STORES the shard's
// key, tagged with an 8-bit epoch, into every peer's mailbox over NVLink (peer-mapped memory, one 8-byte
// relaxed system-scope store per peer: data and arrival in one word).  Warp 0 of EVERY CTA polls this GPU's own
// mailbox row (lane p watches rank p's word) until every word carries the epoch and takes the max — there is no
// second publish hop.  The key embeds the GLOBAL position, so max == the canonical arg-best across ranks.  Rows of the
// mailbox ring are reused 4 barriers later; epochs continue across launches, so nothing is ever reset.
// Wait until every CTA of the grid has arrived at `slot`, then return the final key.  The arrivals are release-adds
// that follow each CTA's atomicMax; the poll is an ACQUIRE load of the counter, so once it reads gridDim.x every
// CTA's max is ordered before the load of `best` that follows (PTX memory model: release/acquire on `arrived`) —
// the guarantee does not lean on how a 128-bit load happens to be served by an L2 sector.
__device__ __forceinline__ unsigned long long slot_poll(OpSync* slot) {
  uint32_t cnt;
  do {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(cnt) : "l"(&slot->arrived) : "memory");
  } while (cnt < gridDim.x);
  unsigned long long best;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(best) : "l"(&slot->best) : "memory");
  return best;
}

#define XGPU_SPIN_LIMIT (1u << 24)       // ~10-20 s of polling before a CTA gives up on a peer (counters[6] = 1)
__device__ __forceinline__ unsigned long long slot_argbest(const KernelArgs& ka, OpSync* slot, uint32_t b,
                                                           unsigned long long warp_key,
                                                           unsigned long long* s_red, unsigned long long* s_win) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  if (lane == 0) s_red[warp] = warp_key;
  __syncthreads();
  if (warp == 0) {
    unsigned long long v = lane < nwarps ? s_red[lane] : 0ull;
    v = warp_max_u64(v);
    if (ka.world <= 1) {
      if (lane == 0) {
        if (v) atomicMax(&slot->best, v);
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&slot->arrived) : "memory");
        *s_win = slot_poll(slot);
      }
    } else {
      const unsigned long long epoch = (unsigned long long)((ka.epoch_base + b + 1u) & 255u);
      const uint32_t row = (ka.epoch_base + b) & 3u;
      if (lane == 0) {
        if (v) atomicMax(&slot->best, v);
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&slot->arrived) : "memory");
        if (blockIdx.x == 0) {      // the rank's leader: wait for this GPU's CTAs, then post the shard's key to every rank
          const unsigned long long best = slot_poll(slot);
          const unsigned long long mine = (best & ~255ull) | epoch;
          for (int p = 1; p <= ka.world; p++) {     // peers first, own mailbox last
            const int q = (ka.rank + p) % ka.world;
            asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(ka.mbox_peer[q] + row * 8 + ka.rank), "l"(mine) : "memory");
          }
        }
      }
      // every CTA polls this GPU's own mailbox row (lane p watches rank p's word): no second publish hop
      const unsigned long long* my_box = ka.mbox_peer[ka.rank] + row * 8;
      unsigned long long w = epoch;
      uint32_t spins = 0;
      bool ok;
      do {
        if (lane < ka.world) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(my_box + lane) : "memory");
        ok = __all_sync(0xffffffffu, (w & 255ull) == epoch);
      } while (!ok && ++spins < XGPU_SPIN_LIMIT);
      if (!ok) { if (lane == 0) ka.counters[6] = 1u; w = 0; }
      w = lane < ka.world ? (w & ~255ull) : 0ull;
      w = warp_max_u64(w);
      if (lane == 0) *s_win = w;
    }
  }
  __syncthreads();
  return *s_win;
}

// The same slot barrier in two halves (one GPU only), so that work whose result is needed only AFTER the winner is
// known — the winner-side full evaluation — runs while the arrivals travel: slot_arrive folds the CTA's key in and
// signals, slot_wait polls.  Every thread calls both; nothing else may touch s_red / s_win in between.
__device__ __forceinline__ void slot_arrive(OpSync* slot, unsigned long long warp_key, unsigned long long* s_red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  if (lane == 0) s_red[warp] = warp_key;
  __syncthreads();
  if (warp == 0) {
    unsigned long long v = lane < nwarps ? s_red[lane] : 0ull;
    v = warp_max_u64(v);
    if (lane == 0) {
      if (v) atomicMax(&slot->best, v);
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&slot->arrived) : "memory");
    }
  }
}
// max of the warp keys slot_arrive left in s_red (every lane reads all of them: broadcast loads, no shuffles)
__device__ __forceinline__ unsigned long long cta_best(const unsigned long long* s_red) {
  const int nwarps = blockDim.x >> 5;
  unsigned long long m = 0;
  for (int w = 0; w < nwarps; w++) {
    const unsigned long long v = s_red[w];
    m = v > m ? v : m;
  }
  return m;
}
__device__ __forceinline__ unsigned long long slot_wait(OpSync* slot, unsigned long long* s_win) {
  if (threadIdx.x == 0) *s_win = slot_poll(slot);
  __syncthreads();
  return *s_win;
}

// The multi-GPU barrier in the same two halves.  xgpu_arrive: every CTA folds its key into the local slot and
// signals; the rank's leader (CTA 0, lane 0 of warp 0) then waits for this GPU's CTAs and posts the shard's key to
// every rank's mailbox.  xgpu_wait: warp 0 of every CTA polls this GPU's own mailbox row.  Between the two the
// groups that fit run their full evaluation, so it overlaps the NVLink round trip instead of preceding it (only
// the leader's own warp 0 does that work after its post, not beside it).
__device__ __forceinline__ void xgpu_arrive(const KernelArgs& ka, OpSync* slot, uint32_t b,
                                            unsigned long long warp_key, unsigned long long* s_red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  if (lane == 0) s_red[warp] = warp_key;
  __syncthreads();
  if (warp == 0) {
    unsigned long long v = lane < nwarps ? s_red[lane] : 0ull;
    v = warp_max_u64(v);
    if (lane == 0) {
      if (v) atomicMax(&slot->best, v);
      asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&slot->arrived) : "memory");
      if (blockIdx.x == 0) {
        const unsigned long long epoch = (unsigned long long)((ka.epoch_base + b + 1u) & 255u);
        const uint32_t row = (ka.epoch_base + b) & 3u;
        const unsigned long long best = slot_poll(slot);
        const unsigned long long mine = (best & ~255ull) | epoch;
        for (int p = 1; p <= ka.world; p++) {     // peers first, own mailbox last
          const int q = (ka.rank + p) % ka.world;
          asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(ka.mbox_peer[q] + row * 8 + ka.rank), "l"(mine) : "memory");
        }
      }
    }
    __syncwarp();
  }
}
__device__ __forceinline__ unsigned long long xgpu_wait(const KernelArgs& ka, uint32_t b, unsigned long long* s_win) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == 0) {
    const unsigned long long epoch = (unsigned long long)((ka.epoch_base + b + 1u) & 255u);
    const uint32_t row = (ka.epoch_base + b) & 3u;
    const unsigned long long* my_box = ka.mbox_peer[ka.rank] + row * 8;
    unsigned long long w = epoch;
    uint32_t spins = 0;
    bool ok;
    do {
      if (lane < ka.world) asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w) : "l"(my_box + lane) : "memory");
      ok = __all_sync(0xffffffffu, (w & 255ull) == epoch);
    } while (!ok && ++spins < XGPU_SPIN_LIMIT);
    if (!ok) { if (lane == 0) ka.counters[6] = 1u; w = 0; }
    w = lane < ka.world ? (w & ~255ull) : 0ull;
    w = warp_max_u64(w);
    if (lane == 0) *s_win = w;
  }
  __syncthreads();
  return *s_win;
}

// ---- request staging: op records (160 B) are prefetched with cp.async into a ring of HS_PF shared-memory
// buffers, HS_PF-1 ops ahead: a RELEASE op is only a few hundred cycles long, shorter than the L2 latency of the
// next record, so a distance of one op stalled every short op on its successor's fetch.  Every op position issues
// exactly one commit_group (an empty one past the end of the stream), so "all but the HS_PF-2 newest groups are
// complete" means "op r has landed".
#define HS_PF 4
__device__ __forceinline__ void prefetch_op(const DevPod* src, DevPod* dst, bool valid) {
  if (valid && threadIdx.x < sizeof(DevPod) / 16) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(reinterpret_cast<char*>(dst) + threadIdx.x * 16);
    const char* g = reinterpret_cast<const char*>(src) + threadIdx.x * 16;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(g) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void prefetch_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void prefetch_wait_oldest() { asm volatile("cp.async.wait_group %0;" ::"n"(HS_PF - 2) : "memory"); }


#define FULL 0xffffffffu

// wall-clock stamp of a decision (ns, %globaltimer): rides in DevDecision::pad so the host can report when each
// decision of a batch became final (bench.py's latency object)
__device__ __forceinline__ void stamp_decision(DevDecision& dec) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  dec.pad[0] = (uint32_t)t; dec.pad[1] = (uint32_t)(t >> 32);
}

// GENERAL = false compiles the loop without the general evaluator (hs_eval.cuh): the host launches that instantiation
// for ranges of ops that are all common-shape Filters and releases — fewer registers, no local-memory frame, and
// more co-resident CTAs, i.e. a larger register-resident inventory.
// The full evaluation of a warp's candidate nodes while the barrier is in flight (whole warp: fast_fit is a collective).
// Resident mode: the group's own node, already in D.  Streaming mode: each group's best node of the sweep, loaded again
// (nothing commits between the sweep and the barrier, so the copy is current); D and my_flags are scratch there.
__device__ __forceinline__ void deferred_eval(const KernelArgs& ka, bool resident, int gg, int my_node, int glane,
                                              uint32_t& my_flags, DevRegs& D, const DevPod& pod, FastEval& fe) {
  int node = gg;
  if (!resident) {
    node = my_node >= 0 ? my_node : 0;
    my_flags = ka.node_flags[node];
    load_dev(ka, node, glane, D);
  }
  fast_fit(D, glane, (uint32_t)node, my_flags, pod, pod.e0, ka, fe);
}

template <bool GENERAL>
__device__ __forceinline__ void decide_loop(const KernelArgs& ka) {
  __shared__ __align__(16) DevPod s_op[HS_PF];
  __shared__ __align__(16) DevEntry s_ent[HS_MAX_E];
  __shared__ unsigned long long s_red[32];
  __shared__ unsigned long long s_win;
  __shared__ unsigned long long s_best[2];    // streaming sweep: best fitting key this CTA has found for op r, in [r & 1]
  __shared__ WinShared s_ws;

  if (threadIdx.x == 0) s_best[0] = s_best[1] = 0ull;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int glane = lane & 7;                 // device slot owned by this lane
  const int gbase = lane & 24;                // first lane of this node's quarter warp
  const uint32_t gmask = 0xffu << gbase;
  const int gg = gtid >> 3;                   // node-owner group id
  const int ngroups = (gridDim.x * blockDim.x) >> 3;
  const bool resident = ka.n_nodes <= ngroups && !ka.no_resident;  // every group owns at most one node: it stays in registers

  DevRegs D;                   // the owned device (resident mode), kept current by this group's own commits
  D.used = D.mem = D.cores = D.count = D.totmem = D.totcore = 0;
  D.packed = D.dflags = 0;
  uint32_t my_flags = 0;
  int my_pos = -1;
  if (resident && gg < ka.n_nodes) {
    my_flags = ka.node_flags[gg];
    load_dev(ka, gg, glane, D);
    if (my_flags & NF_ALIVE) my_pos = ka.cand_pos ? ka.cand_pos[gg] : gg;
  }
  // ComputeDefaultScore of the owned node only changes when its group commits/releases on it: cached
  float my_pre = fast_node_score(D);
  bool dirty = false;          // usage of the owned node changed: refresh my_pre at the next convergent point
  bool reload = false;         // the general path committed through global memory: re-read the slot

  DevAlloc la[GENERAL ? HS_MAX_ALLOC : 1];   // general path (lane 0 of the group): picks of the best owned node for the current op
  uint32_t fail_count = 0;     // no-fit decisions so far: identical in every thread
  if (ka.chain_in != nullptr) {
    // a launch queued behind its predecessor: it only runs if that one got as far as this one's first op (a launch
    // that filled the fail buffer stops early; the host drains the buffer and submits the rest again)
    if (__ldcg(ka.chain_in) != (uint32_t)ka.op_begin) return;     // the same word for every thread of the grid: a uniform exit
    fail_count = __ldcg(ka.chain_in + 1);
  }
  uint32_t barriers = 0;       // arg-best exchanges so far in this launch: identical in every thread (and rank)
  int r = ka.op_begin;
  int ring_base = r;           // op whose record sits in s_op[0]: the prefetch ring restarts after every window
  int no_window_at = -1;       // op that a window left to the one-op path (it found no node: FailedNodes reasons)
  for (int i = 0; i < HS_PF - 1; i++) prefetch_op(ka.pods + r + i, &s_op[i], r + i < ka.n_ops);

  const bool timing = ka.timing != nullptr && gtid == 0;   // diagnostic phase clocks (HS_TIMING=1), block 0 only
  long long tc = timing ? clock64() : 0;
#define TICK(slot) do { if (timing) { const long long t_ = clock64(); ka.timing[slot] += (unsigned long long)(t_ - tc); tc = t_; } } while (0)

  for (; r < ka.n_ops; r++) {
    const int buf = (r - ring_base) & (HS_PF - 1);
    prefetch_wait_oldest();
    __syncthreads();             // op r is staged; every thread is done with op r-1's buffer and with s_ent
    if (threadIdx.x == 0) s_best[(r + 1) & 1] = 0ull;   // op r-1's bound: nobody reads it any more, op r+1 starts from 0
    prefetch_op(ka.pods + r + HS_PF - 1, &s_op[(buf + HS_PF - 1) & (HS_PF - 1)], r + HS_PF - 1 < ka.n_ops);
    const DevPod& pod = s_op[buf];
    // a window works on HBM only: the stale register copies it leaves behind are re-read when a one-op decision needs them
    const bool will_window = ka.w_skey != nullptr && pod.win_len >= 2 && r != no_window_at;
    const bool do_reload = reload && !will_window;
    if (resident && __any_sync(FULL, dirty || do_reload)) {   // warp-uniform: refresh what a commit/release changed
      if (do_reload && gg < ka.n_nodes) load_dev(ka, gg, glane, D);
      my_pre = fast_node_score(D);
      if ((dirty || do_reload) && glane == 0 && gg < ka.n_nodes) ka.node_tot[gg].y = __float_as_int(my_pre);
      dirty = false;
      if (do_reload) reload = false;
    }
    TICK(0);

    if (will_window) {
      // ---- a window: this op and the next win_len-1 decided with one sweep and one barrier (hs_window.cuh)
      int why = 0;
      const int wlen = min(min((int)pod.win_len, HS_WIN), ka.n_ops - r);
      const int n_exec = run_window(ka, r, wlen, s_ws, &why);
      if (why == 1) no_window_at = r + n_exec;
      if (n_exec > 0) {
        reload = resident;       // the window committed through HBM: the register copies are stale
        dirty = false;
        prefetch_wait();
        __syncthreads();
        r += n_exec;
        ring_base = r;
        for (int i = 0; i < HS_PF - 1; i++) prefetch_op(ka.pods + r + i, &s_op[i], r + i < ka.n_ops);
        r--;                     // the loop's own increment
        TICK(7);
        continue;
      }
      // the window's first op found no node: nothing was changed, the one-op path below writes the reasons — on
      // current registers: a reload that an earlier window left pending was skipped above
      if (resident && __any_sync(FULL, reload)) {     // warp-uniform: fast_node_score is a collective
        if (reload && gg < ka.n_nodes) load_dev(ka, gg, glane, D);
        my_pre = fast_node_score(D);
        reload = false;
      }
    }

    if (pod.op == OP_RELEASE) {
      // podManager.DelPod + usage release (scheduler.go:226-253): the owner subtracts what the commit added.
      // An in-batch reference is resolved by each group's lane 0; only the group that wrote that decision is
      // guaranteed to see it, and only that group can match the ownership test.
      int node = pod.rel_node;
      const int4* rows;
      if (node >= 0) rows = ka.rel_pool + pod.rel_off;
      else {
        node = ka.decisions[pod.rel_ref].node;
        if (ka.decisions[pod.rel_ref].status != 0) node = -1;
        node = __shfl_sync(FULL, node, gbase);
        rows = ka.commit_delta + (size_t)pod.rel_ref * HS_D;
      }
      if (node >= 0 && node < ka.n_nodes && (node % ngroups) == gg) {
        int4 delta;
        if (pod.rel_node >= 0) {          // known pod: the rows came with the op record (no second fetch)
          const RelDelta rd = reinterpret_cast<const RelDelta*>(&pod.e0)[glane];
          delta = make_int4((int)rd.slots, rd.mem, (int)rd.cores, 0);
          if (pod.flags & PF_REL_MIG) delta.w = rows[glane].w;     // MIG bookkeeping rides in the pool rows only
        } else {
          delta = rows[glane];
        }
        if (delta.x | delta.y | delta.z | delta.w) {
          int4* p = ka.dyn + (size_t)node * HS_D + glane;
          if (resident) {                 // the registers are the current copy: store, do not read-modify-write
            D.used -= delta.x; D.mem -= delta.y; D.cores -= delta.z;
            if (delta.w) D.dflags = mig_apply(D.dflags, delta.w, -1);
            *p = make_int4(D.used, D.mem, D.cores, (int)D.dflags);
          } else {
            int4 v = *p;
            v.x -= delta.x; v.y -= delta.y; v.z -= delta.z;
            if (delta.w) v.w = (int)mig_apply((uint32_t)v.w, delta.w, -1);
            *p = v;
          }
        }
        dirty = true;
        if (!resident) { __syncwarp(gmask); if (glane == 0) refresh_node_score(ka, node); }
        if (ka.quota != nullptr && (pod.q_idx[0] >= 0 || pod.q_idx[1] >= 0) && glane == 0)
          quota_apply(ka, pod, node, rows, -1);               // quotaManager.RmUsage, scheduler.go:250-252
      }
      if (ka.quota != nullptr && (pod.q_idx[0] >= 0 || pod.q_idx[1] >= 0)) {   // usage of a quota'd namespace changed: publish it
        __threadfence();
        slot_argbest(ka, ka.sync2 + r, 0u, 0ull, s_red, &s_win);
        __threadfence();
      }
      if (gtid == 0) { ka.decisions[r].status = 2; ka.decisions[r].n_alloc = 0; }
      TICK(1);
      continue;
    }
    if (pod.op != OP_FILTER) {
      if (gtid == 0) { ka.decisions[r].status = 2; ka.decisions[r].node = -1; ka.decisions[r].n_alloc = 0; }
      continue;
    }

    const bool spread = pod.flags & PF_NODE_SPREAD;
    const bool dry = pod.flags & PF_DRYRUN;       // filterSimulation (scheduler.go:1156-1199): decide, do not commit
    const bool has_quota = GENERAL && ka.quota != nullptr && (pod.q_idx[0] >= 0 || pod.q_idx[1] >= 0);
    const bool fast = !GENERAL || pod.fast != 0;
    if (!GENERAL && pod.fast == 0) {          // the host never sends such an op to this instantiation
      if (gtid == 0) { ka.decisions[r].status = -1; ka.decisions[r].node = -1; ka.decisions[r].n_alloc = 0; }
      continue;
    }
    // Without numa-bind the VERDICT needs no score/sort (fast_fit_lite).  The full evaluation that tells WHICH devices
    // runs while the barrier's arrivals (and, across GPUs, the mailbox round trip) are in flight (`deferred`), in the
    // warp that holds its CTA's best key — on the register-resident node, or in streaming mode on that warp's best
    // nodes loaded again — so the winner commits straight after the barrier.
    const bool lite = fast && !hs_verdict_needs_order(pod);
    const bool deferred = lite;
    bool evaluated = false;      // warp-uniform: this warp ran the deferred evaluation of its groups' best nodes
    const DevEntry* ent = &pod.e0;
    if (pod.n_entries > 1) {     // general path with several entries: fetch them from the pool
      const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + pod.entry_off);
      uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
      for (int i = threadIdx.x; i < pod.n_entries * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
      __syncthreads();
      ent = s_ent;
    }

    // ---- sweep: every group evaluates the node(s) it owns --------------------------------------
    unsigned long long my_key = 0;   // same value in all 8 lanes of a group
    unsigned long long t_key = 0;    // general path: the best key among the nodes THIS thread evaluated
    int my_node = -1;
    float my_score = 0.0f;
    uint32_t my_nalloc = 0;
    FastEval fe;
    fe.cand = 0; fe.q = 0; fe.slot_of = 0; fe.memreq = 0; fe.cores = 0; fe.net_topo = false;
    const bool net_scored = fast && hs_is_net_scored(pod);   // Ascend910*: ScoreNode depends on the picks (fast_ascend_select)
    int a_row = -1, a_np = 0;
    float a_term = 0.0f;
    if (fast) {
      if (resident) {
        // whole warp, collectives inside; without numa-bind the verdict needs no score/sort (fast_fit_lite)
        const bool fit = lite ? fast_fit_lite(D, glane, gbase, (uint32_t)gg, my_flags, pod, pod.e0, ka)
                              : fast_fit(D, glane, (uint32_t)gg, my_flags, pod, pod.e0, ka, fe);
        if (net_scored) a_np = fast_ascend_select(D, glane, gbase, fit, pod.e0, fe, &a_row, &a_term);   // request-level: warp-uniform
        if (fit && my_pos >= 0) {
          my_score = net_scored ? __fadd_rn(my_pre, a_term) : my_pre;
          my_key = make_key(my_score, spread, (uint32_t)(ka.pos_base + my_pos));
          my_node = gg;
        }
      } else {
        // streaming: the shard does not fit the grid's registers.  Software pipeline as in hs_sweep_kernel: the next
        // node's three 16 B loads are in flight while this node is evaluated; the node score is the cached
        // ComputeDefaultScore in node_tot (every commit/release in this mode refreshes it before the next op).
        // That score — hence the node's key — does not depend on the request, so a node whose key is below the best
        // FITTING key this CTA has found so far cannot win and its verdict is skipped (s_best; the arg-best only needs
        // the maximum, and the maximum is never skipped).  Ascend910* keys carry a request-dependent term: no skipping.
        // Only where a group walks three nodes or more: measured on B200, C5 on one GPU (7 nodes per group) 88.9 k ->
        // 116.4 k decisions/s, C4 (1.4 nodes per group) 140.9 k -> 133.2 k.
        volatile unsigned long long* sb = &s_best[r & 1];
        const bool prune = !net_scored && ka.n_nodes > 2 * ngroups;
        DevRegs n;
        FastEval f2;
        // The group walks its nodes from the HIGHEST index down: equal scores are decided by the position, the higher one
        // wins, so on an inventory of many equal nodes the first node a group looks at already bounds all its others.
        int node = gg < ka.n_nodes ? gg + ((ka.n_nodes - 1 - gg) / ngroups) * ngroups : -1;
        int4 nt = make_int4(0, 0, 0, 0), a = nt, b = nt;
        if (node >= 0) {
          nt = ka.node_tot[node];
          a = ka.dyn[(size_t)node * HS_D + glane];
          b = __ldg(ka.stat + (size_t)node * HS_D + glane);
        }
        while (node >= 0) {   // n_nodes % 32 == 0: warp-uniform trip count
          const int next = node - ngroups;
          int4 nt2 = make_int4(0, 0, 0, 0), a2 = nt2, b2 = nt2;
          if (next >= 0) {
            nt2 = ka.node_tot[next];
            a2 = ka.dyn[(size_t)next * HS_D + glane];
            b2 = __ldg(ka.stat + (size_t)next * HS_D + glane);
          }
          const uint32_t nf = (uint32_t)nt.x;
          const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
          n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
          n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
          bool need = true;
          if (prune) {
            const unsigned long long bound = *sb;
            need = pos >= 0 && make_key(__int_as_float(nt.y), spread, (uint32_t)(ka.pos_base + pos)) > (bound > my_key ? bound : my_key);
          }
          if (__any_sync(FULL, need)) {      // warp-uniform: the verdict is a whole-warp collective
            const bool fit = (lite ? fast_fit_lite(n, glane, gbase, (uint32_t)node, nf, pod, pod.e0, ka)
                                   : fast_fit(n, glane, (uint32_t)node, nf, pod, pod.e0, ka, f2)) && pos >= 0;
            float term2 = 0.0f;
            if (net_scored) { int row2; fast_ascend_select(n, glane, gbase, fit, pod.e0, f2, &row2, &term2); }
            if (fit) {
              const float sc = net_scored ? __fadd_rn(__int_as_float(nt.y), term2) : __int_as_float(nt.y);
              const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
              if (key > my_key) {
                my_key = key; my_node = node; my_score = sc;
                // a plain store, not an atomic: a lost update only leaves the bound lower than it could be, and every
                // value ever stored is the key of a node that fits this op
                if (glane == 0 && prune && key > *sb) *sb = key;
              }
            }
          }
          node = next; nt = nt2; a = a2; b = b2;
        }
      }
    } else if constexpr (GENERAL) {
      // general evaluator: one thread per node.  Resident mode: lane 0 of the group that holds the node.  Streaming
      // mode: the group's node list {gg, gg + ngroups, …} is dealt to its 8 lanes, so a node is always evaluated — and,
      // when it wins, committed — by the same thread, and the 20 000-node configs use every lane instead of one in 8.
      t_key = 0ull;
      if (glane == 0 || !resident) {
        NodeWork w;
        EvalOut out;
        DevAlloc tmp[HS_MAX_ALLOC];
        const int first = resident ? gg : gg + glane * ngroups, stride = resident ? ngroups : ngroups * HS_D;
        for (int node = first; node < ka.n_nodes; node += stride) {
          const uint32_t nf = ka.node_flags[node];
          if (!(nf & NF_ALIVE)) continue;
          const int pos = ka.cand_pos ? ka.cand_pos[node] : node;
          if (pos < 0) continue;
          load_node(ka, node, nf, w);
          out.allocs = tmp; out.n_alloc = 0;
          float sc;
          if (dev_score_node<EV_RECORD>(w, pod, ent, (uint32_t)node, nf, ka, out, &sc)) {
            const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
            if (key > t_key) {
              t_key = key; my_node = node; my_score = sc; my_nalloc = out.n_alloc;
              for (uint32_t i = 0; i < out.n_alloc; i++) la[i] = tmp[i];
            }
          }
        }
      }
      // the group's key = the best of its lanes (my_key stays uniform within the group; my_node etc. stay the thread's own)
      my_key = t_key;
#pragma unroll
      for (int o = 1; o < HS_D; o <<= 1) {
        const unsigned long long t = __shfl_xor_sync(FULL, my_key, o);
        my_key = t > my_key ? t : my_key;
      }
    }
    TICK(2);

    // ---- arg-best: warp shuffle, one atomicMax per CTA, ONE grid barrier -----------------------
    const unsigned long long bk = warp_max_u64(my_key);
    TICK(3);
    unsigned long long win;
    if (ka.world <= 1) {
      slot_arrive(ka.sync + r, bk, s_red);
      if (ka.trace != nullptr && threadIdx.x == 0 && r - ka.op_begin < ka.trace_ops) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        ka.trace[(size_t)(r - ka.op_begin) * ka.trace_stride + blockIdx.x] = t_;
      }
      // Only the warp that holds the CTA's best key can hold the winner: the others skip the full evaluation and
      // leave the issue slots to it (s_red holds every warp's key since slot_arrive's __syncthreads).
      if (deferred && bk != 0ull && bk == cta_best(s_red)) {  // warp-uniform: fast_fit is a whole-warp collective
        evaluated = true;
        deferred_eval(ka, resident, gg, my_node, glane, my_flags, D, pod, fe);
      }
      // (choosing the devices here too, in every group that fits, was measured 8 % SLOWER than leaving fast_pick to
      // the one winner: 5 000 groups pay for it, one benefits)
      win = slot_wait(ka.sync + r, &s_win);
      if (ka.trace != nullptr && gtid == 0 && r - ka.op_begin < ka.trace_ops) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        ka.trace[(size_t)(r - ka.op_begin) * ka.trace_stride + ka.trace_stride - 1] = t_;
      }
    } else if (deferred) {
      xgpu_arrive(ka, ka.sync + r, barriers, bk, s_red);
      if (bk != 0ull && bk == cta_best(s_red)) {
        evaluated = true;
        deferred_eval(ka, resident, gg, my_node, glane, my_flags, D, pod, fe);
      }
      win = xgpu_wait(ka, barriers, &s_win);
    } else {
      win = slot_argbest(ka, ka.sync + r, barriers, bk, s_red, &s_win);
    }
    barriers++;
    TICK(4);

    if (KEY_VALID(win)) {
      const bool mine = my_key == win;   // uniform within the group
      if (ka.world > 1 && gtid == 0) {
        const int lp = KEY_POS(win) - ka.pos_base;
        if (lp < 0 || lp >= ka.n_nodes) {     // the winner lives on another rank's shard
          DevDecision dec;
          dec.status = 3; dec.node = KEY_POS(win); dec.alloc_off = 0; dec.n_alloc = 0; dec.fail_slot = 0xffffffffu;
          uint32_t u = (uint32_t)(win >> 32);
          if (spread) u = ~u;
          u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
          dec.score = __uint_as_float(u);
          stamp_decision(dec);
          ka.decisions[r] = dec;
        }
      }
      if (fast) {
        if (!resident && !evaluated && __any_sync(FULL, mine)) {
          // the sweep only established THAT the node fits (or ran on streamed registers): the winner's warp
          // now runs the full evaluation to learn which devices (collectives → the whole warp takes part)
          if (!resident) {
            const int node = my_node >= 0 ? my_node : 0;
            my_flags = ka.node_flags[node];
            load_dev(ka, node, glane, D);
          }
          const bool wfit = fast_fit(D, glane, (uint32_t)(my_node >= 0 ? my_node : 0), my_flags, pod, pod.e0, ka, fe);
          if (net_scored) a_np = fast_ascend_select(D, glane, gbase, wfit, pod.e0, fe, &a_row, &a_term);
        }
        if (mine) {
          // ---- commit: each lane of the winning group commits its own device ------------------
          int my_row;
          int np;
          if (net_scored) { my_row = a_row; np = a_np; }
          else np = fast_pick(glane, (uint32_t)my_node, my_flags, pod, pod.e0, ka, fe, &my_row);
          const uint32_t off = pod.alloc_base;      // rows of this op: a prefix sum the host made, no cursor atomic
          int4 delta = make_int4(0, 0, 0, 0);
          if (my_row >= 0) {
            delta = make_int4(1, fe.memreq, fe.cores, 0);
            DevAlloc a;
            a.entry = 0; a.dev = (uint8_t)glane; a.usedcores = (int16_t)fe.cores; a.usedmem = fe.memreq;
            if (off + (uint32_t)my_row < ka.alloc_cap) ka.allocs[off + my_row] = a;
            if (!dry) {
              int4* p = ka.dyn + (size_t)my_node * HS_D + glane;
              if (resident) {             // the registers are the current copy: store, do not read-modify-write
                D.used += 1; D.mem += fe.memreq; D.cores += fe.cores;
                *p = make_int4(D.used, D.mem, D.cores, (int)D.dflags);
              } else {
                int4 v = *p;
                v.x += 1; v.y += fe.memreq; v.z += fe.cores;
                *p = v;
              }
            }
          }
          ka.commit_delta[(size_t)r * HS_D + glane] = delta;
          if (glane == 0) {
            if (ka.trace != nullptr && r - ka.op_begin < ka.trace_ops)
              ka.trace[(size_t)(r - ka.op_begin) * ka.trace_stride + ka.trace_stride - 2] = (unsigned long long)blockIdx.x + 1ull;
            DevDecision dec;
            dec.status = 0; dec.node = my_node; dec.score = my_score; dec.alloc_off = off; dec.n_alloc = (uint32_t)np;
            dec.fail_slot = 0xffffffffu; stamp_decision(dec);
            ka.decisions[r] = dec;
          }
          if (!dry) {
            dirty = true;
            if (!resident) { __syncwarp(gmask); if (glane == 0) refresh_node_score(ka, my_node); }
          }
        }
      } else if (GENERAL && mine) {
        if constexpr (GENERAL) {
        if (t_key == win) {          // the one thread that evaluated the winning node
          int4 delta[HS_D];
          collapse_rows(ent, la, my_nalloc, delta);
          const uint32_t off = pod.alloc_base;
          for (uint32_t i = 0; i < my_nalloc && off + i < ka.alloc_cap; i++) ka.allocs[off + i] = la[i];
          const int sg_cls = pod.sg_class & 3;
          if (sg_cls == 1 || sg_cls == 2)          // metax sGPU: the pod joins the PodInfos of every device it received (scheduler.go:786)
            for (uint32_t i = 0; i < my_nalloc; i++)
              if (ent[AL_ENTRY(la[i])].flags & EF_SGPU) delta[AL_DEV(la[i])].w = sg_cls << 2;
          if (!dry) {
            store_usage(ka, my_node, delta, +1);
            if (sg_cls != 1)                          // addJitteryQos (metax/sdevice.go:558-577): a QoS switch is now in flight on these devices
              for (uint32_t i = 0; i < my_nalloc; i++) {
                if (!(ent[AL_ENTRY(la[i])].flags & EF_SGPU)) continue;
                const uint32_t expected = la[i].usedcores == 100 ? 0u : ((uint32_t)pod.sg_qos & 7u);
                if ((uint32_t)AL_MIG_CAND(la[i]) == expected) continue;
                int4* p = ka.dyn + (size_t)my_node * HS_D + AL_DEV(la[i]);
                p->w = (int)(((uint32_t)p->w & 0x0fffffffu) | (1u << 28) | (expected << 29));
              }
            refresh_node_score(ka, my_node);
            quota_apply(ka, pod, my_node, delta, +1);      // quotaManager.AddUsage, scheduler.go:1136-1138
          }
          for (int d = 0; d < HS_D; d++) ka.commit_delta[(size_t)r * HS_D + d] = delta[d];
          DevDecision dec;
          dec.status = 0; dec.node = my_node; dec.score = my_score; dec.alloc_off = off; dec.n_alloc = my_nalloc;
          dec.fail_slot = 0xffffffffu; stamp_decision(dec);
          ka.decisions[r] = dec;
        }
        __syncwarp(gmask);       // lane 0's stores to the slot rows precede the group's reload
        reload = true;
        }
      }
      if (has_quota && !dry) {   // the winner changed a namespace's usage: every CTA must see it before the next sweep
        __threadfence();
        slot_argbest(ka, ka.sync2 + r, 0u, 0ull, s_red, &s_win);
        __threadfence();
      }
    } else {
      // ---- no node fits: per-node reasons for FailedNodes (scheduler.go:1109-1114) ------------
      if (fail_count >= ka.fail_cap) break;  // fail buffer full: the host drains it and resumes at op r
      if (fast) {
        DevFail* fb = ka.fail + (size_t)fail_count * ka.n_nodes;
        DevFail nc;
        nc.kind_cards = FK_NOT_CANDIDATE; nc.c0 = nc.c1 = nc.pad = 0;
        if (resident) {
          const DevFail f = fast_explain(D, glane, (uint32_t)gg, my_flags, pod, pod.e0, ka);   // whole warp
          if (glane == 0 && gg < ka.n_nodes) fb[gg] = my_pos >= 0 ? f : nc;
        } else {
          DevRegs n;
          for (int node = gg; node < ka.n_nodes; node += ngroups) {   // warp-uniform trip count
            const uint32_t nf = ka.node_flags[node];
            const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
            load_dev(ka, node, glane, n);
            const DevFail f = fast_explain(n, glane, (uint32_t)node, nf, pod, pod.e0, ka);
            if (glane == 0) fb[node] = pos >= 0 ? f : nc;
          }
        }
      } else if (GENERAL && (glane == 0 || !resident)) {
        if constexpr (GENERAL) {
        DevFail* fb = ka.fail + (size_t)fail_count * ka.n_nodes;
        NodeWork w;
        EvalOut out;
        const int first = resident ? gg : gg + glane * ngroups, stride = resident ? ngroups : ngroups * HS_D;
        for (int node = first; node < ka.n_nodes; node += stride) {
          const uint32_t nf = ka.node_flags[node];
          DevFail f;
          f.kind_cards = FK_NOT_CANDIDATE; f.c0 = f.c1 = f.pad = 0;
          const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
          if (pos >= 0) {
            load_node(ka, node, nf, w);
            out.fkind = FK_NONE; out.fcards = 0;
            float sc;
            dev_score_node<EV_EXPLAIN>(w, pod, ent, (uint32_t)node, nf, ka, out, &sc);
            pack_fail_node(out, f, node);
          }
          fb[node] = f;
        }
        }
      }
      if (gtid == 0) {
        DevDecision dec;
        dec.status = 1; dec.node = -1; dec.score = 0.0f; dec.alloc_off = 0; dec.n_alloc = 0;
        dec.fail_slot = fail_count; stamp_decision(dec);
        ka.decisions[r] = dec;
      }
      fail_count++;
    }
    TICK(6);
  }
  prefetch_wait();
  if (resident && __any_sync(FULL, dirty)) {   // publish the score of nodes changed by the last one-op commits (a pending
    my_pre = fast_node_score(D);               // reload means HBM is already current: windows and the general path store it)
    if (dirty && glane == 0 && gg < ka.n_nodes) ka.node_tot[gg].y = __float_as_int(my_pre);
  }
  if (gtid == 0) {
    ka.counters[2] = (uint32_t)r; ka.counters[1] = fail_count; ka.counters[5] = barriers;
    if (ka.chain_out != nullptr) { ka.chain_out[0] = (uint32_t)r; ka.chain_out[1] = fail_count; }
  }
}

extern "C" __global__ void __launch_bounds__(512, 1)
hs_decide_kernel(const KernelArgs ka) { decide_loop<true>(ka); }

// the same loop without the general evaluator (launched when a range holds only common-shape Filters and releases)
extern "C" __global__ void __launch_bounds__(256, 3)
hs_decide_fast_kernel(const KernelArgs ka) { decide_loop<false>(ka); }

// Sweep-only kernel: evaluates ONE request against the whole inventory and reduces to the arg-best key
// — no commit.  It is the inventory scan measured for the HBM roofline at inventories larger than L2
//  and the per-rank step of the sharded path.  Common-shape requests stream the node-major
// SoA with one lane per device: a warp reads 4 nodes = 512 contiguous bytes of each array per step
// with 128-bit loads, 256 B per node in total.
// MODE 0: general evaluator (one thread per node); 1: lane-per-device with sort+scan (numa-bind); 2: lane-per-device,
// order-free verdict.  Separate instantiations keep the streaming kernels' register footprint small.
template <int MODE>
__global__ void __launch_bounds__(256, (MODE == 0 ? 2 : (MODE == 1 ? 4 : 5)))
hs_sweep_kernel(const KernelArgs ka, int op) {
  __shared__ __align__(16) DevPod s_pod;
  __shared__ __align__(16) DevEntry s_ent[HS_MAX_E];
  __shared__ unsigned long long s_red[8];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const DevPod& pod = s_pod;
  const DevEntry* ent = &s_pod.e0;
  if (pod.n_entries > 1) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + pod.entry_off);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
    for (int i = threadIdx.x; i < pod.n_entries * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    ent = s_ent;
  }
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const int glane = lane & 7, gbase = lane & 24;
  const int gg = gtid >> 3;
  const int ngroups = (gridDim.x * blockDim.x) >> 3;
  const bool spread = pod.flags & PF_NODE_SPREAD;
  unsigned long long my_key = 0;
  if (MODE != 0) {
    constexpr bool lite = MODE == 2;
    DevRegs n;
    FastEval fe;
    const LiteReq lq = lite_req(pod, pod.e0);   // the request's invariants, in registers for the whole walk
    // software pipeline: the next node's three loads are in flight while this node is evaluated
    int node = gg;
    int4 nt = make_int4(0, 0, 0, 0), a = nt, b = nt;
    if (node < ka.n_nodes) {
      nt = __ldg(ka.node_tot + node);
      a = ka.dyn[(size_t)node * HS_D + glane];
      b = __ldg(ka.stat + (size_t)node * HS_D + glane);
    }
    while (node < ka.n_nodes) {   // n_nodes % 32 == 0: warp-uniform trip count
      const int next = node + ngroups;
      int4 nt2 = make_int4(0, 0, 0, 0), a2 = nt2, b2 = nt2;
      if (next < ka.n_nodes) {
        nt2 = __ldg(ka.node_tot + next);
        a2 = ka.dyn[(size_t)next * HS_D + glane];
        b2 = __ldg(ka.stat + (size_t)next * HS_D + glane);
      }
      const uint32_t nf = (uint32_t)nt.x;
      const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
      n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
      n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
      const bool fit = (lite ? fast_fit_lite_q(n, glane, gbase, (uint32_t)node, nf, lq, pod.e0, ka)
                             : fast_fit(n, glane, (uint32_t)node, nf, pod, pod.e0, ka, fe)) && pos >= 0;
      float term = 0.0f;
      if (!lite && hs_is_net_scored(pod)) { int row_; fast_ascend_select(n, glane, gbase, fit, pod.e0, fe, &row_, &term); }
      {
        const float sc = __fadd_rn(__int_as_float(nt.y), term);   // ComputeDefaultScore, maintained by the commits (+ Ascend910 ScoreNode)
        if (fit) {
          const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
          my_key = key > my_key ? key : my_key;
        }
      }
      node = next; nt = nt2; a = a2; b = b2;
    }
  } else if (MODE == 0 && glane == 0) {
    NodeWork w;
    EvalOut out;
    for (int node = gg; node < ka.n_nodes; node += ngroups) {
      const uint32_t nf = ka.node_flags[node];
      if (!(nf & NF_ALIVE)) continue;
      const int pos = ka.cand_pos ? ka.cand_pos[node] : node;
      if (pos < 0) continue;
      load_node(ka, node, nf, w);
      float sc;
      if (dev_score_node<EV_SCORE>(w, pod, ent, (uint32_t)node, nf, ka, out, &sc)) {
        const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
        my_key = key > my_key ? key : my_key;
      }
    }
  }
  unsigned long long bk = warp_max_u64(my_key);
  if (lane == 0) s_red[threadIdx.x >> 5] = bk;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) bk = s_red[i] > bk ? s_red[i] : bk;
    if (bk) atomicMax(&ka.sync[op].best, bk);
  }
}

// ---- the same scan with the inventory staged through shared memory by the TMA (cp.async.bulk) ----------------------
// A CTA's 32 node groups walk 32 CONSECUTIVE nodes per step, i.e. one contiguous 4 KiB run of `dyn`, one of `stat` and
// 512 B of `node_tot`: three bulk copies into a ring of SW_ST stages, completion counted on one mbarrier per stage.  The
// loads of the next SW_ST - 1 steps are in flight while a step is evaluated, and they cost no registers — the register
// pipeline above holds ONE step ahead, which leaves the scan short of bytes in flight.  Only this GPU's own
// memory is read, and no copy is multicast.
constexpr int SW_ST = 4;
struct __align__(128) SweepStage { int4 dyn[256]; int4 stat[256]; int4 tot[32]; };
static_assert(sizeof(SweepStage) == 8704, "stage layout");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* b, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "SW_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra SW_DONE;\n\t"
      "bra SW_WAIT;\n\t"
      "SW_DONE:\n\t"
      "}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, unsigned long long* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(256, (MODE == 1 ? 4 : 5))
hs_sweep_tma_kernel(const KernelArgs ka, int op) {
  static_assert(MODE == 1 || MODE == 2, "lane-per-device modes only");
  __shared__ SweepStage ring[SW_ST];
  __shared__ __align__(8) unsigned long long full[SW_ST];
  __shared__ __align__(16) DevPod s_pod;
  __shared__ unsigned long long s_red[8];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  if (threadIdx.x == 0) {
    for (int s_ = 0; s_ < SW_ST; s_++) mbar_init(&full[s_], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const DevPod& pod = s_pod;
  const int lane = threadIdx.x & 31;
  const int glane = lane & 7, gbase = lane & 24;
  const bool spread = pod.flags & PF_NODE_SPREAD;
  const int tile0 = (int)blockIdx.x * 32, tstride = (int)gridDim.x * 32;        // n_nodes % 32 == 0: whole tiles only
  const int T = tile0 < ka.n_nodes ? (ka.n_nodes - tile0 + tstride - 1) / tstride : 0;
  auto issue = [&](int k) {
    SweepStage& st = ring[k % SW_ST];
    unsigned long long* b = &full[k % SW_ST];
    const size_t node0 = (size_t)tile0 + (size_t)k * tstride;
    mbar_expect_tx(b, (uint32_t)sizeof(SweepStage));
    bulk_g2s(st.dyn, ka.dyn + node0 * HS_D, sizeof st.dyn, b);
    bulk_g2s(st.stat, ka.stat + node0 * HS_D, sizeof st.stat, b);
    bulk_g2s(st.tot, ka.node_tot + node0, sizeof st.tot, b);
  };
  if (threadIdx.x == 0)
    for (int k = 0; k < T && k < SW_ST; k++) issue(k);
  constexpr bool lite = MODE == 2;
  const LiteReq lq = lite_req(pod, pod.e0);
  unsigned long long my_key = 0;
  DevRegs n;
  FastEval fe;
  // (Measured and not kept: per-stage "empty" mbarriers instead of the CTA barrier below, the walk unrolled over the ring —
  // 0.62 of peak against 0.68; the producer thread's wait stalls its whole warp and the spinning costs issue slots.)
  for (int k = 0; k < T; k++) {
    const SweepStage& st = ring[k % SW_ST];
    mbar_wait(&full[k % SW_ST], (uint32_t)(k / SW_ST) & 1u);
    const int4 a = st.dyn[threadIdx.x], b = st.stat[threadIdx.x], nt = st.tot[threadIdx.x >> 3];
    const int node = tile0 + k * tstride + (int)(threadIdx.x >> 3);
    const uint32_t nf = (uint32_t)nt.x;
    const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
    n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
    n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
    const bool fit = (lite ? fast_fit_lite_q(n, glane, gbase, (uint32_t)node, nf, lq, pod.e0, ka)
                           : fast_fit(n, glane, (uint32_t)node, nf, pod, pod.e0, ka, fe)) && pos >= 0;
    float term = 0.0f;
    if (!lite && hs_is_net_scored(pod)) { int row_; fast_ascend_select(n, glane, gbase, fit, pod.e0, fe, &row_, &term); }
    if (fit) {
      const float sc = __fadd_rn(__int_as_float(nt.y), term);
      const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
      my_key = key > my_key ? key : my_key;
    }
    __syncthreads();                       // every thread has read stage k % SW_ST: it can be filled again
    if (threadIdx.x == 0 && k + SW_ST < T) issue(k + SW_ST);
  }
  unsigned long long bk = warp_max_u64(my_key);
  if (lane == 0) s_red[threadIdx.x >> 5] = bk;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) bk = s_red[i] > bk ? s_red[i] : bk;
    if (bk) atomicMax(&ka.sync[op].best, bk);
  }
}

// ---- sharded path (one process per GPU, nodes partitioned in contiguous blocks) -----------------
// Per op the host enqueues, with no host synchronisation in between:
//   hs_sweep_kernel            → this rank's arg-best key in sync[op].best
//   ncclAllReduce(MAX, uint64) → the global winner in place (the key embeds the GLOBAL position, so max
//                                implements the canonical tie-break across ranks)
//   hs_shard_decide_kernel     → the rank whose shard holds the position commits; the others record "remote"
//   hs_shard_explain_kernel    → only does work when no rank fits: this shard's FailedNodes reasons
// RELEASE ops need only the decide kernel.
extern "C" __global__ void __launch_bounds__(32)
hs_shard_decide_kernel(const KernelArgs ka, int op) {
  __shared__ __align__(16) DevPod s_pod;
  __shared__ __align__(16) DevEntry s_ent[HS_MAX_E];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const DevEntry* ent = &s_pod.e0;
  if (s_pod.op == OP_FILTER && s_pod.n_entries > 1) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + s_pod.entry_off);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
    for (int i = threadIdx.x; i < s_pod.n_entries * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    ent = s_ent;
  }
  if (threadIdx.x != 0) return;
  DevDecision dec;
  dec.status = 2; dec.node = -1; dec.score = 0; dec.alloc_off = 0; dec.n_alloc = 0; dec.fail_slot = 0xffffffffu;
  dec.pad[0] = dec.pad[1] = 0;
  if (s_pod.op == OP_RELEASE) {
    int node = s_pod.rel_node;
    const int4* rows;
    if (node >= 0) rows = ka.rel_pool + s_pod.rel_off;
    else {
      node = ka.decisions[s_pod.rel_ref].status == 0 ? ka.decisions[s_pod.rel_ref].node : -1;   // only the owner rank committed it
      rows = ka.commit_delta + (size_t)s_pod.rel_ref * HS_D;
    }
    if (node >= 0 && node < ka.n_nodes) {
      int4 delta[HS_D];
      for (int d = 0; d < HS_D; d++) delta[d] = rows[d];
      store_usage(ka, node, delta, -1);
      refresh_node_score(ka, node);
    }
    ka.decisions[op] = dec;
    return;
  }
  if (s_pod.op != OP_FILTER) { ka.decisions[op] = dec; return; }
  const unsigned long long win = ka.sync[op].best;   // merged across ranks
  if (!KEY_VALID(win)) {
    dec.status = 1;
    dec.fail_slot = atomicAdd(ka.counters + 1, 1u);  // this shard's slot in the fail buffer (may exceed fail_cap: not recorded)
    ka.decisions[op] = dec;
    return;
  }
  const int pos = KEY_POS(win) - ka.pos_base;
  int node = -1;
  if (pos >= 0 && pos < ka.n_nodes) {
    if (!ka.cand_pos) node = pos;
    else for (int n = 0; n < ka.n_nodes; n++) if (ka.cand_pos[n] == pos) { node = n; break; }
  }
  dec.status = 3;                                    // fits on another rank's shard
  dec.node = KEY_POS(win);                           // global position
  {
    uint32_t u = (uint32_t)(win >> 32);
    if (s_pod.flags & PF_NODE_SPREAD) u = ~u;
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    dec.score = __uint_as_float(u);
  }
  if (node >= 0 && (ka.node_flags[node] & NF_ALIVE)) {
    const uint32_t nf = ka.node_flags[node];
    NodeWork w;
    load_node(ka, node, nf, w);
    DevAlloc la[HS_MAX_ALLOC];
    EvalOut out;
    out.allocs = la; out.n_alloc = 0;
    float sc;
    if (dev_score_node<EV_RECORD>(w, s_pod, ent, (uint32_t)node, nf, ka, out, &sc)) {
      const uint32_t off = atomicAdd(ka.counters + 0, out.n_alloc);
      for (uint32_t i = 0; i < out.n_alloc && off + i < ka.alloc_cap; i++) ka.allocs[off + i] = la[i];
      int4 delta[HS_D];
      collapse_rows(ent, la, out.n_alloc, delta);
      if (!(s_pod.flags & PF_DRYRUN)) {
        store_usage(ka, node, delta, +1);
        refresh_node_score(ka, node);
      }
      for (int d = 0; d < HS_D; d++) ka.commit_delta[(size_t)op * HS_D + d] = delta[d];
      dec.status = 0; dec.node = node; dec.score = sc; dec.alloc_off = off; dec.n_alloc = out.n_alloc;
    } else {
      dec.status = -1;                               // the sweep and the commit disagree: must not happen
    }
  }
  ka.decisions[op] = dec;
}

extern "C" __global__ void __launch_bounds__(256)
hs_shard_explain_kernel(const KernelArgs ka, int op) {
  if (ka.decisions[op].status != 1) return;          // written by hs_shard_decide_kernel earlier on this stream
  const uint32_t slot = ka.decisions[op].fail_slot;
  if (slot >= ka.fail_cap) return;
  __shared__ __align__(16) DevPod s_pod;
  __shared__ __align__(16) DevEntry s_ent[HS_MAX_E];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const DevEntry* ent = &s_pod.e0;
  if (s_pod.n_entries > 1) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + s_pod.entry_off);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
    for (int i = threadIdx.x; i < s_pod.n_entries * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    ent = s_ent;
  }
  DevFail* fb = ka.fail + (size_t)slot * ka.n_nodes;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int glane = threadIdx.x & 7;
  const int gg = gtid >> 3, ngroups = (gridDim.x * blockDim.x) >> 3;
  DevFail nc;
  nc.kind_cards = FK_NOT_CANDIDATE; nc.c0 = nc.c1 = nc.pad = 0;
  if (s_pod.fast) {
    DevRegs n;
    for (int node = gg; node < ka.n_nodes; node += ngroups) {   // warp-uniform trip count (n_nodes % 32 == 0)
      const uint32_t nf = ka.node_flags[node];
      const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
      load_dev(ka, node, glane, n);
      const DevFail f = fast_explain(n, glane, (uint32_t)node, nf, s_pod, s_pod.e0, ka);
      if (glane == 0) fb[node] = pos >= 0 ? f : nc;
    }
  } else if (glane == 0) {
    NodeWork w;
    EvalOut out;
    for (int node = gg; node < ka.n_nodes; node += ngroups) {
      const uint32_t nf = ka.node_flags[node];
      DevFail f = nc;
      const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
      if (pos >= 0) {
        load_node(ka, node, nf, w);
        out.fkind = FK_NONE; out.fcards = 0;
        float sc;
        dev_score_node<EV_EXPLAIN>(w, s_pod, ent, (uint32_t)node, nf, ka, out, &sc);
        pack_fail(out, f);
      }
      fb[node] = f;
    }
  }
}

// Self-check of the branch-free divide (hs_fast.cuh) against the IEEE divide, over the operand space the
// path produces: numerators/denominators that are int32 counters.  Counts mismatching bit patterns.
extern "C" __global__ void hs_divcheck_kernel(unsigned long long n, unsigned long long seed, unsigned long long* bad) {
  unsigned long long local = 0;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    unsigned long long x = (i + seed) * 0x9E3779B97F4A7C15ull;
    x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
    int num, den;
    if (i < (1ull << 24)) { num = (int)(i & 4095u); den = (int)(i >> 12) + 1; }        // exhaustive small grid
    else {
      const int sh_n = (int)((x >> 58) & 31u), sh_d = (int)((x >> 53) & 31u);
      num = (int)((x & 0x7fffffffu) >> sh_n);
      den = (int)(((x >> 22) & 0x7fffffffu) >> sh_d);
      if (den == 0) den = 1;
    }
    const float a = idiv_rn(num, den);
    const float b = __fdiv_rn((float)num, (float)den);
    if (__float_as_uint(a) != __float_as_uint(b)) local++;
  }
  if (local) atomicAdd(bad, local);
}
extern "C" long long hs_launch_divcheck(unsigned long long n, unsigned long long seed) {
  unsigned long long* d = nullptr;
  unsigned long long hbad = 0;
  if (cudaMalloc(&d, sizeof *d) != cudaSuccess) return -1;
  cudaMemset(d, 0, sizeof *d);
  hs_divcheck_kernel<<<1184, 256>>>(n, seed, d);
  if (cudaMemcpy(&hbad, d, sizeof hbad, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaFree(d); return -1; }
  cudaFree(d);
  return (long long)hbad;
}

// ---- host-callable launch wrappers (kept here so hs_engine.cu stays plain C++) ---------------
extern "C" int hs_launch_decide(const KernelArgs* ka, int grid, int block, cudaStream_t stream, int fast_only) {
  void* args[] = {(void*)ka};
  return (int)cudaLaunchCooperativeKernel(fast_only ? (const void*)hs_decide_fast_kernel : (const void*)hs_decide_kernel, dim3(grid), dim3(block), args, 0, stream);
}
extern "C" int hs_decide_max_blocks(int block, int device, int fast_only) {
  int per_sm = 0, sms = 0;
  cudaError_t e = fast_only ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hs_decide_fast_kernel, block, 0)
                            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hs_decide_kernel, block, 0);
  if (e != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
  return per_sm * sms;
}
// mode: 0 general, 1 lane-per-device full, 2 lane-per-device order-free (see hs_sweep_kernel)
static int sweep_tma_enabled() {      // HS_SWEEP_TMA=0 keeps the register-pipelined scan (A/B measurements)
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("HS_SWEEP_TMA");
    v = (e && e[0] == '0') ? 0 : 1;
    cudaFuncSetAttribute(hs_sweep_tma_kernel<1>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    cudaFuncSetAttribute(hs_sweep_tma_kernel<2>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  }
  return v;
}
extern "C" int hs_launch_sweep(const KernelArgs* ka, int op, int mode, int grid, int block, cudaStream_t stream) {
  if (mode != 0 && block == 256 && sweep_tma_enabled()) {
    if (mode == 2) hs_sweep_tma_kernel<2><<<grid, block, 0, stream>>>(*ka, op);
    else hs_sweep_tma_kernel<1><<<grid, block, 0, stream>>>(*ka, op);
    return (int)cudaGetLastError();
  }
  if (mode == 2) hs_sweep_kernel<2><<<grid, block, 0, stream>>>(*ka, op);
  else if (mode == 1) hs_sweep_kernel<1><<<grid, block, 0, stream>>>(*ka, op);
  else hs_sweep_kernel<0><<<grid, block, 0, stream>>>(*ka, op);
  return (int)cudaGetLastError();
}
extern "C" int hs_sweep_blocks_per_sm(int mode) { return mode == 0 ? 2 : (mode == 1 ? 4 : 5); }
extern "C" int hs_launch_shard_decide(const KernelArgs* ka, int op, cudaStream_t stream) {
  hs_shard_decide_kernel<<<1, 32, 0, stream>>>(*ka, op);
  return (int)cudaGetLastError();
}
extern "C" int hs_launch_shard_explain(const KernelArgs* ka, int op, int grid, cudaStream_t stream) {
  hs_shard_explain_kernel<<<grid, 256, 0, stream>>>(*ka, op);
  return (int)cudaGetLastError();
}
