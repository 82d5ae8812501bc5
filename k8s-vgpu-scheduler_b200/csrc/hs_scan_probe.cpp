// hs_scan_probe — measurement tool: the inventory scan (hs_sweep_kernel) in isolation on inventories
// from L2-resident to several times L2 (SURVEY §8d "honest HBM measurement").  Builds N synthetic nodes
// through the public C ABI, then times the sweep of one request with CUDA events (hs_sweep_probe).
// Prints one JSON object per (size, request shape).  Usage: hs_scan_probe [nodes ...]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/hami_sched.h"

int main(int argc, char** argv) {
  std::vector<long> sizes;
  for (int i = 1; i < argc; i++) sizes.push_back(atol(argv[i]));
  if (sizes.empty()) sizes = {5000, 100000, 1000000, 4000000};
  for (long n : sizes) {
    hs_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.cuda_device = -1;
    cfg.reserve_nodes = (uint32_t)n;
    hs_handle* h = hs_create(&cfg);
    if (!h) { fprintf(stderr, "hs_create: %s\n", hs_last_error(nullptr)); return 1; }
    char name[64], ids[8][64];
    for (long i = 0; i < n; i++) {
      hs_device_info d[8];
      for (int k = 0; k < 8; k++) {
        snprintf(ids[k], sizeof ids[k], "GPU-%07ld-%d", i, k);
        d[k].id = ids[k]; d[k].index = (uint32_t)k; d[k].count = 10; d[k].devmem = 184320; d[k].devcore = 100;
        d[k].type = "NVIDIA B200"; d[k].numa = k / 4; d[k].mode = "hami-core"; d[k].health = 1; d[k].vendor = "NVIDIA";
        d[k].network_id = -1;
      }
      snprintf(name, sizeof name, "node-%07ld", i);
      hs_node_info ni;
      memset(&ni, 0, sizeof ni);
      ni.name = name; ni.devices = d; ni.n_devices = 8; ni.hami_vnpu_core = -1;
      if (hs_node_upsert(h, &ni) != HS_OK) { fprintf(stderr, "upsert: %s\n", hs_last_error(h)); return 1; }
    }
    for (int shape = 0; shape < 2; shape++) {
      hs_container_request rq = {"NVIDIA", "NVIDIA", 2, 0, 50, 30};
      hs_container c = {&rq, 1};
      hs_kv annos[2] = {{"hami.io/node-scheduler-policy", "spread"}, {"nvidia.com/numa-bind", "true"}};
      hs_pod pod;
      memset(&pod, 0, sizeof pod);
      pod.uid = "probe"; pod.ns = "default"; pod.name = "probe";
      pod.annotations = annos; pod.n_annotations = shape == 0 ? 1 : 2;
      pod.containers = &c; pod.n_containers = 1;
      double ms = 0; int32_t idx = -1; float score = 0;
      const uint32_t iters = n >= 1000000 ? 20 : 200;
      int rc = hs_sweep_probe(h, &pod, iters, &ms, &idx, &score);
      if (rc < 0) { fprintf(stderr, "probe: %s\n", hs_last_error(h)); return 1; }
      const double bytes = (double)((n + 31) / 32 * 32) * 256.0;
      printf("{\"kernel\": \"hs_sweep_kernel\", \"nodes\": %ld, \"request\": \"%s\", \"inventory_mb\": %.2f, \"ms\": %.5f, "
             "\"algorithmic_gbps\": %.1f, \"winner\": %d, \"iters\": %u}\n",
             n, shape == 0 ? "nums=2 mempct=50 spread (order-free fit)" : "same + numa-bind (full sort+scan)",
             bytes / 1e6, ms, bytes / (ms * 1e-3) / 1e9, idx, iters);
      fflush(stdout);
    }
    hs_destroy(h);
  }
  return 0;
}
