// hs_kernels.cu — sm_100a kernels of the HAMi Filter/Score engine.
//
// hs_decide_kernel is ONE persistent cooperative kernel that runs a whole stream
// of scheduling ops in order.  Per FILTER op (reference: Scheduler.Filter,
// pkg/scheduler/scheduler.go:1067-1154):
//   1. the op's request descriptor is staged in shared memory;
//   2. every thread evaluates the node(s) it owns (hs_eval.cuh: ComputeScore, the
//      stable device sort, the vendor Fit scan, ComputeDefaultScore) reading the
//      SoA inventory with coalesced 128-bit loads — or, when every thread owns at
//      most one node, from its registers/local copy loaded once per launch;
//   3. warp-shuffle + shared-memory arg-best, one 64-bit atomicMax per CTA;
//   4. ONE grid barrier;
//   5. the thread that owns the winning node commits the collapsed usage
//      (CollapseInitContainerUsage + AddPod's effect on the next usage rebuild) and
//      writes the decision record.  Ownership of a node is static, so the commit is
//      only ever read back by the thread that wrote it: no second barrier.
// A decision therefore sees every earlier commit, exactly like the reference's
// sequential Filter calls, and costs one barrier instead of a kernel launch.
//
// Integer/compare work on CUDA cores: there is no dense contraction on this path,
// so no tensor cores.
#include <cuda_runtime.h>

#include "hs_eval.cuh"
#include "hs_fast.cuh"



__device__ __forceinline__ void load_node(const KernelArgs& ka, int node, uint32_t nflags, NodeWork& w) {
  const int tile = node / HS_TILE, lane = node % HS_TILE;
  const int4* dy = ka.dyn + (size_t)tile * HS_D * HS_TILE + lane;
  const int4* st = ka.stat + (size_t)tile * HS_D * HS_TILE + lane;
  int4 a[HS_D], b[HS_D];
#pragma unroll
  for (int d = 0; d < HS_D; d++) { a[d] = dy[d * HS_TILE]; b[d] = __ldg(st + d * HS_TILE); }
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    w.used[d] = a[d].x; w.mem[d] = a[d].y; w.cores[d] = a[d].z; w.dflags[d] = (uint32_t)a[d].w;
    w.count[d] = b[d].x; w.totmem[d] = b[d].y; w.totcore[d] = b[d].z; w.packed[d] = (uint32_t)b[d].w;
  }
  w.ndev = NF_NDEV(nflags);
  w.order = 0x76543210u;
}

__device__ __forceinline__ void reload_usage(const KernelArgs& ka, int node, NodeWork& w) {
  const int tile = node / HS_TILE, lane = node % HS_TILE;
  const int4* dy = ka.dyn + (size_t)tile * HS_D * HS_TILE + lane;
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    const int4 a = dy[d * HS_TILE];
    w.used[d] = a.x; w.mem[d] = a.y; w.cores[d] = a.z;
  }
}

// CollapseInitContainerUsage (pkg/device/initContainer.go:36-114) folded into the usage
// the next rebuild would see (scheduler.go:782-785): per device slot {slots, mem, cores}.
__device__ void collapse_rows(const DevEntry* ent, const DevAlloc* al, uint32_t n, int4* delta) {
  for (int d = 0; d < HS_D; d++) delta[d] = make_int4(0, 0, 0, 0);
  // groups are (type key, device); visit each group once at its first row
  for (uint32_t i = 0; i < n; i++) {
    const int key = ent[al[i].entry].key_id, dev = al[i].dev;
    bool first = true;
    for (uint32_t j = 0; j < i; j++)
      if (ent[al[j].entry].key_id == key && al[j].dev == dev) { first = false; break; }
    if (!first) continue;
    int imem = 0, icores = 0, islots = 0, amem = 0, acores = 0, aslots = 0;
    for (uint32_t j = i; j < n; j++) {
      if (ent[al[j].entry].key_id != key || al[j].dev != dev) continue;
      if (ent[al[j].entry].is_init) {
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

__device__ __forceinline__ void store_usage(const KernelArgs& ka, int node, const int4* delta, int sign) {
  const int tile = node / HS_TILE, lane = node % HS_TILE;
  int4* dy = ka.dyn + (size_t)tile * HS_D * HS_TILE + lane;
  for (int d = 0; d < HS_D; d++) {
    if (delta[d].x == 0 && delta[d].y == 0 && delta[d].z == 0) continue;
    int4 v = dy[d * HS_TILE];
    v.x += sign * delta[d].x; v.y += sign * delta[d].y; v.z += sign * delta[d].z;
    dy[d * HS_TILE] = v;
  }
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
  if (out.fkind == FK_HIST) {
    for (int r = 0; r < 8; r++) c0 |= (out.reasons[r] & 15u) << (4 * r);
    for (int r = 8; r < 15; r++) c1 |= (out.reasons[r] & 15u) << (4 * (r - 8));
  }
  f.c0 = c0; f.c1 = c1; f.pad = 0;
}

// ---- grid-wide barrier ---------------------------------------------------------------------------
// One arrival counter in L2, never reset inside a launch: barrier k completes when it reaches
// (k+1) * gridDim.x.  The launch is cooperative, so every CTA is resident.  Thread 0 publishes this
// CTA's arg-best (atomicMax) before its release-add, polls with relaxed loads (no L1 invalidation
// inside the spin), then performs one acquire load.
__device__ __forceinline__ void grid_barrier(uint32_t* arrive, uint32_t expected) {
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(arrive) : "memory");
    uint32_t v;
    do {
      asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(arrive) : "memory");
    } while (v < expected);
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(arrive) : "memory");
  }
  __syncthreads();
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// ---- request staging: the next op's 160 B record is prefetched with cp.async while the current
// op is evaluated (two shared-memory buffers).
__device__ __forceinline__ void prefetch_op(const DevPod* src, DevPod* dst) {
  if (threadIdx.x < sizeof(DevPod) / 16) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(reinterpret_cast<char*>(dst) + threadIdx.x * 16);
    const char* g = reinterpret_cast<const char*>(src) + threadIdx.x * 16;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(g) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void prefetch_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void load_regs(const KernelArgs& ka, int node, NodeRegs& n) {
  const int tile = node / HS_TILE, lane = node % HS_TILE;
  const int4* dy = ka.dyn + (size_t)tile * HS_D * HS_TILE + lane;
  const int4* st = ka.stat + (size_t)tile * HS_D * HS_TILE + lane;
  int4 a[HS_D], b[HS_D];
#pragma unroll
  for (int d = 0; d < HS_D; d++) { a[d] = dy[d * HS_TILE]; b[d] = __ldg(st + d * HS_TILE); }
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    n.used[d] = a[d].x; n.mem[d] = a[d].y; n.cores[d] = a[d].z; n.dflags[d] = (uint32_t)a[d].w;
    n.count[d] = b[d].x; n.totmem[d] = b[d].y; n.totcore[d] = b[d].z; n.packed[d] = (uint32_t)b[d].w;
  }
}

__device__ __forceinline__ void regs_to_work(const NodeRegs& n, uint32_t nflags, NodeWork& w) {
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    w.used[d] = n.used[d]; w.mem[d] = n.mem[d]; w.cores[d] = n.cores[d]; w.dflags[d] = n.dflags[d];
    w.count[d] = n.count[d]; w.totmem[d] = n.totmem[d]; w.totcore[d] = n.totcore[d]; w.packed[d] = n.packed[d];
  }
  w.ndev = NF_NDEV(nflags);
  w.order = 0x76543210u;
}

extern "C" __global__ void __launch_bounds__(128, 1)
hs_decide_kernel(const KernelArgs ka) {
  __shared__ __align__(16) DevPod s_op[2];
  __shared__ __align__(16) DevEntry s_ent[HS_MAX_E];
  __shared__ unsigned long long s_red[4];

  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gsize = gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool resident = ka.n_nodes <= gsize;  // every thread owns at most one node: it stays in registers

  NodeRegs reg;                // the owned node (resident mode): usage kept current by this thread's own commits
  uint32_t my_flags = 0;
  int my_pos = -1;
#pragma unroll
  for (int d = 0; d < HS_D; d++) {
    reg.used[d] = reg.mem[d] = reg.cores[d] = reg.count[d] = reg.totmem[d] = reg.totcore[d] = 0;
    reg.packed[d] = reg.dflags[d] = 0;
  }
  if (resident && gtid < ka.n_nodes) {
    my_flags = ka.node_flags[gtid];
    load_regs(ka, gtid, reg);
    if (my_flags & NF_ALIVE) my_pos = ka.cand_pos ? ka.cand_pos[gtid] : gtid;
  }

  DevAlloc la[HS_MAX_ALLOC];   // general path: picks of this thread's best node for the current op
  uint32_t fail_count = 0;     // no-fit decisions so far: identical in every thread
  uint32_t barriers = 0;       // grid barriers executed in this launch: identical in every thread
  int r = ka.op_begin;
  if (r < ka.n_ops) prefetch_op(ka.pods + r, &s_op[0]);

  for (; r < ka.n_ops; r++) {
    const int buf = (r - ka.op_begin) & 1;
    prefetch_wait();
    __syncthreads();             // op r is staged; every thread is done with the other buffer and with s_ent
    if (r + 1 < ka.n_ops) prefetch_op(ka.pods + r + 1, &s_op[buf ^ 1]);
    const DevPod& pod = s_op[buf];

    if (pod.op == OP_RELEASE) {
      // podManager.DelPod + usage release (scheduler.go:226-253): the owner subtracts what the commit added.
      // An in-batch reference is only trusted by the thread that wrote that decision (its own stores).
      int node = pod.rel_node;
      const int4* rows;
      if (node >= 0) rows = ka.rel_pool + pod.rel_off;
      else {
        node = ka.decisions[pod.rel_ref].node;
        rows = ka.commit_delta + (size_t)pod.rel_ref * HS_D;
        if (ka.decisions[pod.rel_ref].status != 0) node = -1;
      }
      if (node >= 0 && node < ka.n_nodes && (node % gsize) == gtid) {
        int4 delta[HS_D];
#pragma unroll
        for (int d = 0; d < HS_D; d++) delta[d] = rows[d];
        store_usage(ka, node, delta, -1);
        if (resident) {
#pragma unroll
          for (int d = 0; d < HS_D; d++) { reg.used[d] -= delta[d].x; reg.mem[d] -= delta[d].y; reg.cores[d] -= delta[d].z; }
        }
      }
      if (gtid == 0) { ka.decisions[r].status = 2; ka.decisions[r].n_alloc = 0; }
      continue;
    }
    if (pod.op != OP_FILTER) {
      if (gtid == 0) { ka.decisions[r].status = 2; ka.decisions[r].node = -1; ka.decisions[r].n_alloc = 0; }
      continue;
    }

    const bool spread = pod.flags & PF_NODE_SPREAD;
    const bool fast = pod.fast != 0;
    const DevEntry* ent = &pod.e0;
    if (pod.n_entries > 1) {     // general path with several entries: fetch them from the pool
      const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + pod.entry_off);
      uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
      for (int i = threadIdx.x; i < pod.n_entries * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
      __syncthreads();
      ent = s_ent;
    }

    // ---- sweep: evaluate owned nodes --------------------------------------------------------
    unsigned long long my_key = 0;
    int my_node = -1;
    float my_score = 0.0f;
    uint32_t my_nalloc = 0;
    FastEval fe;
    fe.cand = 0; fe.qpos = 0; fe.cores = 0;
    if (fast) {
      if (resident) {
        if (my_pos >= 0 && fast_fit(reg, (uint32_t)gtid, my_flags, pod, pod.e0, ka, fe)) {
          my_score = fast_node_score(reg);
          my_key = make_key(my_score, spread, (uint32_t)(ka.pos_base + my_pos));
          my_node = gtid;
        }
      } else {
        NodeRegs n;
        FastEval f2;
        for (int node = gtid; node < ka.n_nodes; node += gsize) {
          const uint32_t nf = ka.node_flags[node];
          if (!(nf & NF_ALIVE)) continue;
          const int pos = ka.cand_pos ? ka.cand_pos[node] : node;
          if (pos < 0) continue;
          load_regs(ka, node, n);
          if (fast_fit(n, (uint32_t)node, nf, pod, pod.e0, ka, f2)) {
            const float sc = fast_node_score(n);
            const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
            if (key > my_key) { my_key = key; my_node = node; my_score = sc; }
          }
        }
      }
    } else {
      NodeWork w;
      EvalOut out;
      if (resident) {
        if (my_pos >= 0) {
          regs_to_work(reg, my_flags, w);
          out.allocs = la; out.n_alloc = 0;
          float sc;
          if (dev_score_node<EV_RECORD>(w, pod, ent, (uint32_t)gtid, my_flags, ka, out, &sc)) {
            my_key = make_key(sc, spread, (uint32_t)(ka.pos_base + my_pos));
            my_node = gtid; my_score = sc; my_nalloc = out.n_alloc;
          }
        }
      } else {
        DevAlloc tmp[HS_MAX_ALLOC];
        for (int node = gtid; node < ka.n_nodes; node += gsize) {
          const uint32_t nf = ka.node_flags[node];
          if (!(nf & NF_ALIVE)) continue;
          const int pos = ka.cand_pos ? ka.cand_pos[node] : node;
          if (pos < 0) continue;
          load_node(ka, node, nf, w);
          out.allocs = tmp; out.n_alloc = 0;
          float sc;
          if (dev_score_node<EV_RECORD>(w, pod, ent, (uint32_t)node, nf, ka, out, &sc)) {
            const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
            if (key > my_key) {
              my_key = key; my_node = node; my_score = sc; my_nalloc = out.n_alloc;
              for (uint32_t i = 0; i < out.n_alloc; i++) la[i] = tmp[i];
            }
          }
        }
      }
    }

    // ---- arg-best: warp shuffle, one atomicMax per CTA, ONE grid barrier -----------------------
    unsigned long long bk = warp_max_u64(my_key);
    if (blockDim.x > 32) {
      if (lane == 0) s_red[warp] = bk;
      __syncthreads();
      if (threadIdx.x == 0)
        for (int i = 1; i < (int)(blockDim.x >> 5); i++) bk = s_red[i] > bk ? s_red[i] : bk;
    }
    if (threadIdx.x == 0 && bk) atomicMax(ka.best + r, bk);
    barriers++;
    grid_barrier(ka.counters + 3, barriers * gridDim.x);
    const unsigned long long win = ld_relaxed_u64(ka.best + r);

    if (win != 0) {
      if (my_key == win) {
        // ---- commit (owner only) -----------------------------------------------------------
        int4 delta[HS_D];
        if (fast) {
          const uint32_t nf = resident ? my_flags : ka.node_flags[my_node];
          if (!resident) {   // streaming mode: re-derive the winner's picks (reg is free in this mode)
            load_regs(ka, my_node, reg);
            fast_fit(reg, (uint32_t)my_node, nf, pod, pod.e0, ka, fe);
          }
          int picks[HS_D];
          const int np = fast_pick((uint32_t)my_node, nf, pod, pod.e0, ka, fe, picks);
          my_nalloc = (uint32_t)np;
#pragma unroll
          for (int d = 0; d < HS_D; d++) delta[d] = make_int4(0, 0, 0, 0);
          for (int i = 0; i < np; i++) {
            const int s = picks[i];
            int memreq = 0;
#pragma unroll
            for (int d = 0; d < HS_D; d++) if (d == s) memreq = fe.memreq[d];
            la[i].entry = 0; la[i].dev = (uint8_t)s; la[i].usedcores = (int16_t)fe.cores; la[i].usedmem = memreq;
#pragma unroll
            for (int d = 0; d < HS_D; d++) if (d == s) delta[d] = make_int4(1, memreq, fe.cores, 0);
          }
        } else {
          collapse_rows(ent, la, my_nalloc, delta);
        }
        const uint32_t off = atomicAdd(ka.counters + 0, my_nalloc);
        for (uint32_t i = 0; i < my_nalloc && off + i < ka.alloc_cap; i++) ka.allocs[off + i] = la[i];
        store_usage(ka, my_node, delta, +1);
        if (resident) {
#pragma unroll
          for (int d = 0; d < HS_D; d++) { reg.used[d] += delta[d].x; reg.mem[d] += delta[d].y; reg.cores[d] += delta[d].z; }
        }
#pragma unroll
        for (int d = 0; d < HS_D; d++) ka.commit_delta[(size_t)r * HS_D + d] = delta[d];
        DevDecision dec;
        dec.status = 0; dec.node = my_node; dec.score = my_score; dec.alloc_off = off; dec.n_alloc = my_nalloc;
        dec.fail_slot = 0xffffffffu; dec.pad[0] = dec.pad[1] = 0;
        ka.decisions[r] = dec;
      }
    } else {
      // ---- no node fits: per-node reasons for FailedNodes (scheduler.go:1109-1114) ------------
      if (fail_count >= ka.fail_cap) break;  // fail buffer full: the host drains it and resumes at op r
      DevFail* fb = ka.fail + (size_t)fail_count * ka.n_nodes;
      NodeWork w;
      EvalOut out;
      for (int node = gtid; node < ka.n_nodes; node += gsize) {
        const uint32_t nf = ka.node_flags[node];
        DevFail f;
        f.kind_cards = FK_NOT_CANDIDATE; f.c0 = f.c1 = f.pad = 0;
        const int pos = (nf & NF_ALIVE) ? (ka.cand_pos ? ka.cand_pos[node] : node) : -1;
        if (pos >= 0) {
          if (resident) regs_to_work(reg, nf, w);
          else load_node(ka, node, nf, w);
          out.fkind = FK_NONE; out.fcards = 0;
          float sc;
          dev_score_node<EV_EXPLAIN>(w, pod, ent, (uint32_t)node, nf, ka, out, &sc);
          pack_fail(out, f);
        }
        fb[node] = f;
      }
      if (gtid == 0) {
        DevDecision dec;
        dec.status = 1; dec.node = -1; dec.score = 0.0f; dec.alloc_off = 0; dec.n_alloc = 0;
        dec.fail_slot = fail_count; dec.pad[0] = dec.pad[1] = 0;
        ka.decisions[r] = dec;
      }
      fail_count++;
    }
  }
  prefetch_wait();
  if (gtid == 0) { ka.counters[2] = (uint32_t)r; ka.counters[1] = fail_count; }
}

// Sweep-only kernel: evaluates ONE request against the whole inventory and reduces to
// the arg-best key — no commit.  It is the inventory scan measured for the HBM roofline
// at inventories larger than L2 and the per-rank step of the sharded path.
extern "C" __global__ void __launch_bounds__(256)
hs_sweep_kernel(const KernelArgs ka, int op) {
  __shared__ DevPod s_pod;
  __shared__ DevEntry s_ent[HS_MAX_E];
  __shared__ unsigned long long s_red[8];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  {
    const int ne = s_pod.n_entries;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + s_pod.entry_off);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
    for (int i = threadIdx.x; i < ne * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gsize = gridDim.x * blockDim.x;
  const bool spread = s_pod.flags & PF_NODE_SPREAD;
  unsigned long long my_key = 0;
  NodeWork w;
  EvalOut out;
  for (int node = gtid; node < ka.n_nodes; node += gsize) {
    const uint32_t nf = ka.node_flags[node];
    if (!(nf & NF_ALIVE)) continue;
    const int pos = ka.cand_pos ? ka.cand_pos[node] : node;
    if (pos < 0) continue;
    load_node(ka, node, nf, w);
    float sc;
    if (dev_score_node<EV_SCORE>(w, s_pod, s_ent, (uint32_t)node, nf, ka, out, &sc)) {
      const unsigned long long key = make_key(sc, spread, (uint32_t)(ka.pos_base + pos));
      my_key = key > my_key ? key : my_key;
    }
  }
  unsigned long long bk = warp_max_u64(my_key);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = bk;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) bk = s_red[i] > bk ? s_red[i] : bk;
    if (bk) atomicMax(ka.best + op, bk);
  }
}

// Commit / explain for a winner chosen outside the kernel (sharded path: the key was merged
// across ranks; only the rank whose shard holds the position commits).
extern "C" __global__ void __launch_bounds__(32)
hs_commit_kernel(const KernelArgs ka, int op, unsigned long long win) {
  __shared__ DevPod s_pod;
  __shared__ DevEntry s_ent[HS_MAX_E];
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.pods + op);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&s_pod);
    for (int i = threadIdx.x; i < (int)(sizeof(DevPod) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  {
    const int ne = s_pod.n_entries;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(ka.entries + s_pod.entry_off);
    uint32_t* dst = reinterpret_cast<uint32_t*>(s_ent);
    for (int i = threadIdx.x; i < ne * (int)(sizeof(DevEntry) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int pos = (int)((win & 0xffffffffull) - 1ull) - ka.pos_base;
  int node = -1;
  if (!ka.cand_pos) node = pos;
  else for (int n = 0; n < ka.n_nodes; n++) if (ka.cand_pos[n] == pos) { node = n; break; }
  DevDecision dec;
  dec.status = -1; dec.node = node; dec.score = 0; dec.alloc_off = 0; dec.n_alloc = 0; dec.fail_slot = 0xffffffffu;
  dec.pad[0] = dec.pad[1] = 0;
  if (node >= 0 && node < ka.n_nodes) {
    const uint32_t nf = ka.node_flags[node];
    NodeWork w;
    load_node(ka, node, nf, w);
    DevAlloc la[HS_MAX_ALLOC];
    EvalOut out;
    out.allocs = la; out.n_alloc = 0;
    float sc;
    if (dev_score_node<EV_RECORD>(w, s_pod, s_ent, (uint32_t)node, nf, ka, out, &sc)) {
      const uint32_t off = atomicAdd(ka.counters + 0, out.n_alloc);
      for (uint32_t i = 0; i < out.n_alloc && off + i < ka.alloc_cap; i++) ka.allocs[off + i] = la[i];
      int4 delta[HS_D];
      collapse_rows(s_ent, la, out.n_alloc, delta);
      store_usage(ka, node, delta, +1);
      for (int d = 0; d < HS_D; d++) ka.commit_delta[(size_t)op * HS_D + d] = delta[d];
      dec.status = 0; dec.score = sc; dec.alloc_off = off; dec.n_alloc = out.n_alloc;
    }
  }
  ka.decisions[op] = dec;
}

// ---- host-callable launch wrappers (kept here so hs_engine.cu stays plain C++) ---------------
extern "C" int hs_launch_decide(const KernelArgs* ka, int grid, int block, cudaStream_t stream) {
  void* args[] = {(void*)ka};
  return (int)cudaLaunchCooperativeKernel((const void*)hs_decide_kernel, dim3(grid), dim3(block), args, 0, stream);
}
extern "C" int hs_decide_max_blocks(int block, int device) {
  int per_sm = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hs_decide_kernel, block, 0) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
  return per_sm * sms;
}
extern "C" int hs_launch_sweep(const KernelArgs* ka, int op, int grid, int block, cudaStream_t stream) {
  hs_sweep_kernel<<<grid, block, 0, stream>>>(*ka, op);
  return (int)cudaGetLastError();
}
extern "C" int hs_launch_commit(const KernelArgs* ka, int op, unsigned long long win, cudaStream_t stream) {
  hs_commit_kernel<<<1, 32, 0, stream>>>(*ka, op, win);
  return (int)cudaGetLastError();
}
