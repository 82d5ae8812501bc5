// hs_scan_probe — measurement tool: the inventory scan (hs_sweep_kernel) in isolation on inventories
// from L2-resident to several times L2 (SURVEY §8d "honest HBM measurement").  Builds N synthetic nodes
// through the public C ABI, then times the sweep of one request with CUDA events (hs_sweep_probe).
// Prints one JSON object per (size, request shape).
//   hs_scan_probe [nodes ...]               idle, homogeneous B200 inventory (the easy case)
//   hs_scan_probe --realistic [nodes ...]   BASELINE config 4's mix: 50 % NVIDIA (4 models), 20 % Ascend910B3, 15 % MLU,
//                                           15 % DCU, 2 % unhealthy devices, ~70 % of the devices in use with uneven
//                                           memory / core shares; requests with type and UUID selectors, numa-bind
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/hami_sched.h"

static unsigned long long g_rng = 0x9E3779B97F4A7C15ull;
static inline unsigned rnd() { g_rng = g_rng * 6364136223846793005ull + 1442695040888963407ull; return (unsigned)(g_rng >> 33); }

struct Shape { const char* what; int nums; int mempct; int cores; std::vector<hs_kv> annos; };

int main(int argc, char** argv) {
  std::vector<long> sizes;
  bool realistic = false;
  for (int i = 1; i < argc; i++) {
    if (!strcmp(argv[i], "--realistic")) realistic = true;
    else sizes.push_back(atol(argv[i]));
  }
  if (sizes.empty()) sizes = {5000, 100000, 1000000, 4000000};
  static const char* kModels[4] = {"NVIDIA A100-SXM4-80GB", "NVIDIA H100 80GB HBM3", "NVIDIA A100-PCIE-40GB", "NVIDIA L40S"};
  static const int kModelMem[4] = {81920, 81559, 40960, 46068};
  for (long n : sizes) {
    hs_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.cuda_device = -1;
    cfg.reserve_nodes = (uint32_t)n;
    hs_ascend_template tmpl[2] = {{"vir05_1c_16g", 16384}, {"vir10_3c_32g", 32768}};
    hs_ascend_config asc = {"Ascend910B3", 65536, 65536, tmpl, 2};
    cfg.ascend = &asc; cfg.n_ascend = 1;
    hs_handle* h = hs_create(&cfg);
    if (!h) { fprintf(stderr, "hs_create: %s\n", hs_last_error(nullptr)); return 1; }
    char name[64], ids[8][64], uid[64];
    std::string uuid_list;               // a use-gpuuuid list of four NVIDIA devices spread over the inventory
    long used_devices = 0, unhealthy = 0, next_mark = n / 8, picked = 0;
    g_rng = 0x9E3779B97F4A7C15ull;
    for (long i = 0; i < n; i++) {
      hs_device_info d[8];
      memset(d, 0, sizeof d);          // the optional parts (MIG profiles, vendor CustomInfo) stay absent
      const unsigned kind = realistic ? rnd() % 100 : 0;     // <50 NVIDIA, <70 Ascend, <85 MLU, else DCU
      const int model = (int)(rnd() % 4);
      for (int k = 0; k < 8; k++) {
        d[k].index = (uint32_t)k; d[k].numa = k / 4; d[k].network_id = -1; d[k].mode = "";
        d[k].health = (realistic && rnd() % 100 < 2) ? 0 : 1;
        unhealthy += !d[k].health;
        if (!realistic) {
          snprintf(ids[k], sizeof ids[k], "GPU-%07ld-%d", i, k);
          d[k].count = 10; d[k].devmem = 184320; d[k].devcore = 100; d[k].type = "NVIDIA B200"; d[k].mode = "hami-core"; d[k].vendor = "NVIDIA";
        } else if (kind < 50) {
          snprintf(ids[k], sizeof ids[k], "GPU-%07ld-%d", i, k);
          d[k].count = 10; d[k].devmem = kModelMem[model]; d[k].devcore = 100; d[k].type = kModels[model]; d[k].mode = "hami-core"; d[k].vendor = "NVIDIA";
        } else if (kind < 70) {
          snprintf(ids[k], sizeof ids[k], "NPU-%07ld-%d", i, k);
          d[k].count = 20; d[k].devmem = 65536; d[k].devcore = 20; d[k].type = "Ascend910B3"; d[k].vendor = "Ascend910B3"; d[k].network_id = k / 4;
        } else if (kind < 85) {
          snprintf(ids[k], sizeof ids[k], "node-%07ld-cambricon-mlu-%d", i, k);
          d[k].count = 100; d[k].devmem = 49152; d[k].devcore = 100; d[k].type = "MLU"; d[k].vendor = "MLU"; d[k].numa = 0;
        } else {
          snprintf(ids[k], sizeof ids[k], "DCU-%07ld-%d", i, k);
          d[k].count = 4; d[k].devmem = 65536; d[k].devcore = 100; d[k].type = (i & 1) ? "DCU-K100" : "DCU-Z100L"; d[k].vendor = "DCU";
        }
        d[k].id = ids[k];
      }
      snprintf(name, sizeof name, "node-%07ld", i);
      hs_node_info ni;
      memset(&ni, 0, sizeof ni);
      ni.name = name; ni.devices = d; ni.n_devices = 8; ni.hami_vnpu_core = -1;
      if (hs_node_upsert(h, &ni) != HS_OK) { fprintf(stderr, "upsert: %s\n", hs_last_error(h)); return 1; }
      if (realistic) {
        // ~70 % of the devices hold something: 1-3 slots, 5-90 % of the memory, 0-100 % of the cores — every node differs
        snprintf(uid, sizeof uid, "held-%07ld", i);
        for (int k = 0; k < 8; k++) {
          if (rnd() % 100 >= 70) continue;
          const int slots = 1 + (int)(rnd() % 3);
          const int mem = (int)((long long)d[k].devmem * (5 + rnd() % 86) / 100);
          const int cores = (int)((long long)d[k].devcore * (rnd() % 101) / 100);
          if (hs_usage_apply(h, uid, name, ids[k], slots, mem, cores) != HS_OK) { fprintf(stderr, "usage: %s\n", hs_last_error(h)); return 1; }
          used_devices++;
        }
        if (kind < 50 && i >= next_mark && picked < 4) {   // the first NVIDIA node after each quarter mark
          if (!uuid_list.empty()) uuid_list += ",";
          uuid_list += ids[3];
          picked++; next_mark += n / 4;
        }
      }
    }
    std::vector<Shape> shapes;
    shapes.push_back({"nums=2 mempct=50 cores=30 spread (order-free verdict)", 2, 50, 30, {{"hami.io/node-scheduler-policy", "spread"}}});
    if (realistic) {
      shapes.push_back({"same + nvidia.com/use-gputype=A100 (type selector)", 2, 50, 30,
                        {{"hami.io/node-scheduler-policy", "spread"}, {"nvidia.com/use-gputype", "A100"}}});
      shapes.push_back({"nums=1 + nvidia.com/use-gpuuuid (4 UUIDs: per-node list lookup)", 1, 20, 10,
                        {{"hami.io/node-scheduler-policy", "spread"}, {"nvidia.com/use-gpuuuid", uuid_list.c_str()}}});
    }
    shapes.push_back({"nums=2 + numa-bind (per-NUMA verdict)", 2, 50, 30,
                      {{"hami.io/node-scheduler-policy", "spread"}, {"nvidia.com/numa-bind", "true"}}});
    shapes.push_back({"nums=2 + numa-bind under gpu policy mutex (sort + scan per node, MODE 1)", 2, 50, 30,
                      {{"hami.io/node-scheduler-policy", "spread"}, {"nvidia.com/numa-bind", "true"}, {"hami.io/gpu-scheduler-policy", "mutex"}}});
    for (auto& sh : shapes) {
      hs_container_request rq = {"NVIDIA", "NVIDIA", sh.nums, 0, sh.mempct, sh.cores};
      hs_container c = {&rq, 1};
      hs_pod pod;
      memset(&pod, 0, sizeof pod);
      pod.uid = "probe"; pod.ns = "default"; pod.name = "probe";
      pod.annotations = sh.annos.data(); pod.n_annotations = (uint32_t)sh.annos.size();
      pod.containers = &c; pod.n_containers = 1;
      double ms = 0; int32_t idx = -1; float score = 0;
      const uint32_t iters = getenv("HS_PROBE_ITERS") ? (uint32_t)atoi(getenv("HS_PROBE_ITERS")) : (n >= 1000000 ? 20 : 200);   // (a profiler run takes 2)
      int rc = hs_sweep_probe(h, &pod, iters, &ms, &idx, &score);
      if (rc < 0) { fprintf(stderr, "probe: %s\n", hs_last_error(h)); return 1; }
      const double nodes_padded = (double)((n + 31) / 32 * 32);
      const double bytes = nodes_padded * 256.0;              // N x 8 x 32 B
      const char* tma_env = getenv("HS_SWEEP_TMA");
      const char* kname = (tma_env && tma_env[0] == '0') ? "hs_sweep_kernel" : "hs_sweep_tma_kernel";
      printf("{\"kernel\": \"%s\", \"inventory\": \"%s\", \"nodes\": %ld, \"request\": \"%s\", \"inventory_mb\": %.2f, \"ms\": %.5f, "
             "\"algorithmic_gbps\": %.1f, \"with_node_records_gbps\": %.1f, \"winner\": %d, \"fits\": %s, \"iters\": %u, \"devices_in_use\": %ld, \"unhealthy\": %ld}\n",
             kname, realistic ? "realistic (C4 mix, ~70% in use, 2% unhealthy)" : "idle homogeneous B200", n, sh.what, bytes / 1e6, ms,
             bytes / (ms * 1e-3) / 1e9, nodes_padded * 272.0 / (ms * 1e-3) / 1e9, idx, rc == 0 ? "true" : "false", iters, used_devices, unhealthy);
      fflush(stdout);
    }
    hs_destroy(h);
  }
  return 0;
}
