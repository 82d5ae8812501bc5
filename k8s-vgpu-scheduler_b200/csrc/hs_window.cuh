// hs_window.cuh — a WINDOW of up to HS_WIN consecutive ops decided with ONE sweep of the inventory and ONE grid
// barrier, bit-identical to deciding them one at a time.
//
// The reference decides pods strictly one after the other (Scheduler.Filter, pkg/scheduler/scheduler.go:1067-1154:
// every Filter sees the commits of all earlier ones).  Between two consecutive decisions only ONE node changes, so:
//
//   phase A  (every CTA)  the inventory is swept ONCE against the state S0 at the start of the window: per 32-node tile
//            each node's arg-best key (node score + canonical position, the same packed key the one-op path reduces)
//            and one fit bit per FILTER op of the window.  The node score (ComputeDefaultScore, node_policy.go:106-130)
//            does not depend on the request, so the tile is sorted ONCE by key (warp bitonic sort) and each op's
//            candidates are a 32-bit mask in that order: `rmask[tile][op]`, plus the tile's best fitting key
//            `head[op][tile]`.
//   barrier  every CTA arrives; only the resolver CTA waits for the arrivals.
//   phase C  (resolver CTA) walks the ops IN ORDER.  Nodes touched so far in the window (a commit or a release) are
//            TRACKED: a quarter warp holds the node's current state in registers and re-evaluates it exactly for every
//            later op (the same fast_fit / fast_pick / fast_node_score as the one-op path).  Every other node still
//            has its S0 state, so its S0 key and fit bit are still exact: the winner of an op is
//                max( re-evaluated tracked nodes , best S0 key among the UNtracked fitting nodes ),
//            the second term being a scan over the tiles' heads (a tile that holds tracked nodes masks them out of its
//            rmask and reads the next key of its sorted list).  An untracked winner is loaded, tracked, and the op is
//            evaluated again (now exactly, on its registers).  Commits and releases are stored to HBM straight away.
//   publish  the resolver release-stores how many ops it decided; the other CTAs acquire it and move on.
// A window stops early — the remaining ops start a new window or take the one-op path — when an op finds no node
// (the FailedNodes reasons need every node's state at that very point: the one-op path writes them) or when the
// tracker is full.  Nothing is approximated and nothing is speculative in the results: every decision is the exact
// arg-best over the exact state its predecessors left.
#pragma once
#include "hs_fast.cuh"
#include "hs_types.cuh"

#define HS_WHEAD_TILES 256   // inventories up to 8,192 nodes keep the window's heads in shared memory during phase C

struct __align__(16) WinShared {
  DevPod op[HS_WIN];
  unsigned long long head[HS_WIN][HS_WHEAD_TILES];   // phase C (resolver): w_head of this window, when n_tiles fits
  unsigned long long key[2][HS_TILE];
  uint32_t fit[2][HS_TILE];
  int tnode[HS_WSLOTS];
  int trank[HS_WSLOTS];
  uint32_t tbits[HS_WMAX_TILES / 32];
  unsigned long long red[8];
  unsigned long long cand[HS_WIN];
  unsigned long long word;
  int n_tracked;
  int n_new;
};

__device__ __forceinline__ unsigned long long win_block_max(unsigned long long v, WinShared& ws) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
    v = t > v ? t : v;
  }
  if (lane == 0) ws.red[warp] = v;
  __syncthreads();
  unsigned long long b = ws.red[0];
#pragma unroll
  for (int i = 1; i < 8; i++) b = ws.red[i] > b ? ws.red[i] : b;
  __syncthreads();
  return b;
}

// Returns the number of ops decided, [r0, r0 + n); *stop_reason: 0 = the whole window, 1 = op r0+n finds no node
// (one-op path), 2 = tracker full (a new window starts at r0+n).  Every thread of the grid calls it (blockDim 256).
static __device__ int run_window(const KernelArgs& ka, int r0, int wlen, WinShared& ws, int* stop_reason) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int glane = lane & 7, gbase = lane & 24;
  const int grp = tid >> 3;                       // 0..31: node of the tile (phase A) / tracker slot (phase C)
  const int n_tiles = ka.n_nodes / HS_TILE;
  OpSync* slot = ka.sync2 + r0;                   // the window's own barrier slot (sync[r0] stays clean for the one-op path)
  static_assert(HS_WIN == 16 && HS_WSLOTS == 32, "phase C maps ops and tracker slots onto the lanes of one warp");
  const bool wt = ka.timing != nullptr && blockIdx.x == 0 && tid == 0;   // HS_TIMING=1: phase clocks of the resolver, slots 8..11
  long long wc = wt ? clock64() : 0;
#define WTICK(slot_) do { if (wt) { const long long t_ = clock64(); ka.timing[slot_] += (unsigned long long)(t_ - wc); wc = t_; } } while (0)

  // ---- the window's op records
  {
    const int4* src = reinterpret_cast<const int4*>(ka.pods + r0);
    int4* dst = reinterpret_cast<int4*>(ws.op);
    for (int i = tid; i < wlen * (int)(sizeof(DevPod) / 16); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  uint32_t fmask = 0;
  bool spread = false;
  for (int j = 0; j < wlen; j++)
    if (ws.op[j].op == OP_FILTER) { fmask |= 1u << j; spread = ws.op[j].flags & PF_NODE_SPREAD; }

  // ---- phase A: one sweep at S0 for every op of the window
  for (int tile = blockIdx.x, it = 0; tile < n_tiles; tile += gridDim.x, it++) {
    const int node = tile * HS_TILE + grp;
    const int4 nt = ka.node_tot[node];
    const int4 a = ka.dyn[(size_t)node * HS_D + glane];
    const int4 b = __ldg(ka.stat + (size_t)node * HS_D + glane);
    DevRegs n;
    n.used = a.x; n.mem = a.y; n.cores = a.z; n.dflags = (uint32_t)a.w;
    n.count = b.x; n.totmem = b.y; n.totcore = b.z; n.packed = (uint32_t)b.w;
    const uint32_t nf = (uint32_t)nt.x;
    uint32_t fits = 0;
    for (int j = 0; j < wlen; j++) {
      if (!((fmask >> j) & 1u)) continue;
      const DevPod& p = ws.op[j];
      bool fit;
      if (hs_verdict_needs_order(p)) { FastEval fe; fit = fast_fit(n, glane, (uint32_t)node, nf, p, p.e0, ka, fe); }
      else fit = fast_fit_lite(n, glane, gbase, (uint32_t)node, nf, p, p.e0, ka);
      fits |= (fit ? 1u : 0u) << j;
    }
    const bool alive = nf & NF_ALIVE;
    const int buf = it & 1;
    if (glane == 0) {
      ws.key[buf][grp] = alive ? make_key(__int_as_float(nt.y), spread, (uint32_t)(ka.pos_base + node)) : 0ull;
      ws.fit[buf][grp] = alive ? fits : 0u;
    }
    __syncthreads();
    if (warp == 0) {
      // sort the tile best key first (bitonic, descending); the original lane rides in the key's free low bits
      unsigned long long k = ws.key[buf][lane] | (unsigned long long)lane;
      uint32_t f = ws.fit[buf][lane];
#pragma unroll
      for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
          const unsigned long long ok = __shfl_xor_sync(0xffffffffu, k, stride);
          const uint32_t of = __shfl_xor_sync(0xffffffffu, f, stride);
          const bool desc_block = (lane & size) == 0 || size == 32;
          const bool lower = (lane & stride) == 0;
          const bool want_max = lower == desc_block;
          const bool take = want_max ? (ok > k) : (ok < k);
          if (take) { k = ok; f = of; }
        }
      }
      ka.w_rank[(size_t)tile * HS_TILE + (int)(k & 31ull)] = (uint8_t)lane;
      k = (k & ~255ull) | (unsigned long long)lane;            // low bits = rank in the sorted tile
      ka.w_skey[(size_t)tile * HS_TILE + lane] = k;
      uint32_t my_m = 0;
      unsigned long long my_h = 0;
      for (int j = 0; j < wlen; j++) {
        const uint32_t m = __ballot_sync(0xffffffffu, (f >> j) & 1u);
        const unsigned long long h = __shfl_sync(0xffffffffu, k, m ? (__ffs(m) - 1) : 0);
        if (lane == j) { my_m = m; my_h = m ? h : 0ull; }
      }
      if (lane < wlen) {
        ka.w_rmask[(size_t)tile * HS_WIN + lane] = my_m;
        ka.w_head[(size_t)lane * n_tiles + tile] = my_h;
      }
    }
  }

  // ---- barrier: every CTA arrives, only the resolver waits for all of them
  __syncthreads();
  WTICK(8);
  if (tid == 0) {
    __threadfence();
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(&slot->arrived) : "memory");
  }
  if (blockIdx.x != 0) {
    if (tid == 0) {
      unsigned long long w;
      do {
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(&slot->best) : "memory");
      } while (w == 0ull);
      ws.word = w;
    }
    __syncthreads();
    const unsigned long long w = ws.word;
    __syncthreads();
    *stop_reason = (int)(w >> 32);
    return (int)(w & 0xffffffffull) - 1;
  }
  if (tid == 0) {
    uint32_t cnt;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(cnt) : "l"(&slot->arrived) : "memory");
    } while (cnt < gridDim.x);
  }
  for (int i = tid; i < HS_WMAX_TILES / 32; i += blockDim.x) ws.tbits[i] = 0u;
  if (tid < HS_WSLOTS) { ws.tnode[tid] = -1; ws.trank[tid] = 0; }
  if (tid == 0) ws.n_tracked = 0;
  __syncthreads();
  WTICK(9);
  const bool heads_in_smem = n_tiles <= HS_WHEAD_TILES;
  if (heads_in_smem) {
    for (int i = tid; i < wlen * n_tiles; i += blockDim.x) ws.head[i / n_tiles][i % n_tiles] = ka.w_head[i];
    __syncthreads();
  }

  // ---- phase C: the resolver decides the ops in order
  DevRegs T;                       // this quarter warp's tracked node (slot = grp)
  T.used = T.mem = T.cores = T.count = T.totmem = T.totcore = 0;
  T.packed = T.dflags = 0;
  uint32_t tflags = 0;
  float tpre = 0.0f;
  int tnode = -1;
  int n_exec = 0, stop = 0;
  uint32_t late = 0;

  // best S0 key of op j among the untracked fitting nodes of tile t
  auto tile_best = [&](int j, int t, int ntr) -> unsigned long long {
    unsigned long long h = heads_in_smem ? ws.head[j][t] : ka.w_head[(size_t)j * n_tiles + t];
    if ((ws.tbits[t >> 5] >> (t & 31)) & 1u) {
      uint32_t m = ka.w_rmask[(size_t)t * HS_WIN + j];
      for (int i = 0; i < ntr; i++)
        if (ws.tnode[i] / HS_TILE == t) m &= ~(1u << ws.trank[i]);
      h = m ? ka.w_skey[(size_t)t * HS_TILE + (__ffs(m) - 1)] : 0ull;
    }
    return h;
  };
  // the quarter warp of slot `grp` loads its node's current state from HBM
  auto load_slot = [&]() {
    const int node = ws.tnode[grp];
    const int4 nt = ka.node_tot[node];
    const int4 a = ka.dyn[(size_t)node * HS_D + glane];
    const int4 b = __ldg(ka.stat + (size_t)node * HS_D + glane);
    T.used = a.x; T.mem = a.y; T.cores = a.z; T.dflags = (uint32_t)a.w;
    T.count = b.x; T.totmem = b.y; T.totcore = b.z; T.packed = (uint32_t)b.w;
    tflags = (uint32_t)nt.x;
    tpre = __int_as_float(nt.y);                 // kept current by every commit and release
    tnode = node;
  };
  // make `node` a tracked node: the next free quarter warp takes it
  auto track = [&](int node, int rank) {
    const int s = ws.n_tracked;
    if (tid == 0) {
      ws.tnode[s] = node;
      ws.trank[s] = rank >= 0 ? rank : (int)ka.w_rank[node];
      ws.tbits[(node / HS_TILE) >> 5] |= 1u << ((node / HS_TILE) & 31);
    }
    __syncthreads();
    if (grp == s) load_slot();
    if (tid == 0) ws.n_tracked = s + 1;
    __syncthreads();
    return s;
  };
  // track up to 32 candidates at once (one per lane of warp 0: node or -1, rank or -1): duplicates and nodes already
  // tracked drop out, the loads of all new slots are in flight together
  auto track_batch = [&](int node, int rank) {
    const int n0 = ws.n_tracked;
    if (warp == 0) {
      bool want = node >= 0 && node < ka.n_nodes;
      for (int i = 0; i < n0; i++) if (ws.tnode[i] == node) want = false;
      const uint32_t peers = __match_any_sync(0xffffffffu, want ? node : -1 - lane);
      const bool leader = want && (__ffs(peers) - 1) == lane;
      const uint32_t lm = __ballot_sync(0xffffffffu, leader);
      const int idx = n0 + __popc(lm & ((1u << lane) - 1u));
      if (leader && idx < HS_WSLOTS) {
        ws.tnode[idx] = node;
        ws.trank[idx] = rank >= 0 ? rank : (int)ka.w_rank[node];
        atomicOr(&ws.tbits[(node / HS_TILE) >> 5], 1u << ((node / HS_TILE) & 31));
      }
      if (lane == 0) ws.n_new = min(__popc(lm), HS_WSLOTS - n0);
    }
    __syncthreads();
    const int nn = ws.n_new;
    if (grp >= n0 && grp < n0 + nn) load_slot();
    __syncthreads();
    if (tid == 0) ws.n_tracked = n0 + nn;
    __syncthreads();
  };

  // ---- pre-track what the window will touch: the nodes its releases name, and (two rounds) each Filter's best
  // untracked candidate — so that the ops below find their winner already in registers
  for (int round = 0; round < 2; round++) {
    const int ntr = ws.n_tracked;
    for (int j = warp; j < wlen; j += 8) {                     // warp w takes ops w, w+8
      unsigned long long k = 0;
      if ((fmask >> j) & 1u)
        for (int t = lane; t < n_tiles; t += 32) { const unsigned long long h = tile_best(j, t, ntr); k = h > k ? h : k; }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { const unsigned long long t2 = __shfl_xor_sync(0xffffffffu, k, o); k = t2 > k ? t2 : k; }
      if (lane == 0) ws.cand[j] = k;
    }
    __syncthreads();
    int cnode = -1, crank = -1;
    if (warp == 0) {
      if (lane < wlen) {
        const unsigned long long k = ws.cand[lane];
        if (KEY_VALID(k)) { cnode = KEY_POS(k) - ka.pos_base; crank = (int)(k & 31ull); }
      } else if (round == 0 && lane - 16 < wlen && lane >= 16) {
        const DevPod& q = ws.op[lane - 16];
        if (q.op == OP_RELEASE) {
          if (q.rel_node >= 0) cnode = q.rel_node;
          else if (q.rel_ref < r0 && ka.decisions[q.rel_ref].status == 0) cnode = ka.decisions[q.rel_ref].node;
        }
      }
    }
    track_batch(cnode, crank);
  }
  WTICK(10);

  for (int j = 0; j < wlen; j++) {
    const DevPod& p = ws.op[j];
    const int r = r0 + j;
    if (p.op == OP_RELEASE) {
      // podManager.DelPod + usage release (scheduler.go:226-253)
      int node = p.rel_node;
      if (node < 0) node = ka.decisions[p.rel_ref].status == 0 ? ka.decisions[p.rel_ref].node : -1;
      if (node >= 0 && node < ka.n_nodes) {
        int s = -1;
        for (int i = 0; i < ws.n_tracked; i++) if (ws.tnode[i] == node) s = i;
        if (s < 0) {
          if (ws.n_tracked == HS_WSLOTS) { stop = 2; break; }
          s = track(node, -1);
        }
        if (grp == s) {
          int4 delta;
          if (p.rel_node >= 0) {
            const RelDelta rd = reinterpret_cast<const RelDelta*>(&p.e0)[glane];
            delta = make_int4((int)rd.slots, rd.mem, (int)rd.cores, 0);
          } else {
            delta = ka.commit_delta[(size_t)p.rel_ref * HS_D + glane];
          }
          if (delta.x | delta.y | delta.z) {
            T.used -= delta.x; T.mem -= delta.y; T.cores -= delta.z;
            ka.dyn[(size_t)node * HS_D + glane] = make_int4(T.used, T.mem, T.cores, (int)T.dflags);
          }
        }
        if (warp == (s >> 2)) {                  // warp-uniform: the node score follows the usage
          const float sc = fast_node_score(T);
          if (grp == s) { tpre = sc; if (glane == 0) ka.node_tot[node].y = __float_as_int(sc); }
        }
      }
      if (tid == 0) { ka.decisions[r].status = 2; ka.decisions[r].n_alloc = 0; }
      n_exec++;
      WTICK(15);
      continue;
    }
    if (p.op != OP_FILTER) {
      if (tid == 0) { ka.decisions[r].status = 2; ka.decisions[r].node = -1; ka.decisions[r].n_alloc = 0; }
      n_exec++;
      continue;
    }
    FastEval fe;
    int win_slot = -1;
    for (;;) {
      // tracked nodes: exact, on their current registers (whole-CTA call: fast_fit is a warp collective)
      const bool fit = fast_fit(T, glane, (uint32_t)(tnode >= 0 ? tnode : 0), tflags, p, p.e0, ka, fe);
      WTICK(12);
      unsigned long long key = 0;
      if (tnode >= 0 && fit && (tflags & NF_ALIVE) && glane == 0)
        key = make_key(tpre, spread, (uint32_t)(ka.pos_base + tnode)) | 0x80ull | (unsigned long long)grp;
      // untracked nodes: their S0 keys, tile by tile
      const int nt_ = ws.n_tracked;
      for (int t = tid; t < n_tiles; t += blockDim.x) { const unsigned long long h = tile_best(j, t, nt_); key = h > key ? h : key; }
      const unsigned long long best = win_block_max(key, ws);
      WTICK(13);
      if (!KEY_VALID(best)) { stop = 1; break; }
      if (best & 0x80ull) { win_slot = (int)(best & 31ull); break; }
      if (ws.n_tracked == HS_WSLOTS) { stop = 2; break; }
      track(KEY_POS(best) - ka.pos_base, (int)(best & 31ull));   // an untracked node wins: load it, evaluate again
      late++;
    }
    if (stop) break;
    // ---- commit: each lane of the winning quarter warp commits its own device (AddResourceUsage + AddPod's effect)
    if (grp == win_slot) {
      int my_row;
      const int np = fast_pick(glane, (uint32_t)tnode, tflags, p, p.e0, ka, fe, &my_row);
      const uint32_t off = p.alloc_base;
      int4 delta = make_int4(0, 0, 0, 0);
      if (my_row >= 0) {
        delta = make_int4(1, fe.memreq, fe.cores, 0);
        DevAlloc al;
        al.entry = 0; al.dev = (uint8_t)glane; al.usedcores = (int16_t)fe.cores; al.usedmem = fe.memreq;
        if (off + (uint32_t)my_row < ka.alloc_cap) ka.allocs[off + my_row] = al;
        T.used += 1; T.mem += fe.memreq; T.cores += fe.cores;
        ka.dyn[(size_t)tnode * HS_D + glane] = make_int4(T.used, T.mem, T.cores, (int)T.dflags);
      }
      ka.commit_delta[(size_t)r * HS_D + glane] = delta;
      if (glane == 0) {
        DevDecision dec;
        dec.status = 0; dec.node = tnode; dec.score = tpre; dec.alloc_off = off; dec.n_alloc = (uint32_t)np;
        dec.fail_slot = 0xffffffffu;
        unsigned long long tns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
        dec.pad[0] = (uint32_t)tns; dec.pad[1] = (uint32_t)(tns >> 32);
        ka.decisions[r] = dec;
      }
    }
    if (warp == (win_slot >> 2)) {
      const float sc = fast_node_score(T);
      if (grp == win_slot) { tpre = sc; if (glane == 0) ka.node_tot[tnode].y = __float_as_int(sc); }
    }
    n_exec++;
    WTICK(14);
  }

  // ---- publish
  WTICK(11);
  __syncthreads();
  if (tid == 0) {
    if (ka.w_stats) {
      ka.w_stats[0] += 1; ka.w_stats[1] += (uint32_t)n_exec; ka.w_stats[4] += late;
      if (stop == 1) ka.w_stats[2] += 1;
      if (stop == 2) ka.w_stats[3] += 1;
    }
    __threadfence();
    const unsigned long long w = ((unsigned long long)stop << 32) | (unsigned long long)(n_exec + 1);
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(&slot->best), "l"(w) : "memory");
  }
  *stop_reason = stop;
  return n_exec;
}
