import sys, time, ctypes as C
sys.path.insert(0, '/root/repo')
import importlib
eng = importlib.import_module('k8s_vgpu_scheduler_b200.engine')
gen = importlib.import_module('k8s_vgpu_scheduler_b200.gen')
abi = importlib.import_module('k8s_vgpu_scheduler_b200.abi')
cfg, nodes, ops = gen.CONFIGS['c3'](n_nodes=5000, n_pods=9000)
e = eng.Engine(cfg['node_policy'], cfg['gpu_policy'])
for n in nodes: e.node_upsert(n)
e.run_batch(eng.PackedOps(ops[:8000]))
L = e.L
L.hs_debug_host_seconds.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
pods = [a for k, a in ops[8000:] if k == 'filter'][:600]
built = [abi.build_pod(p) for p in pods]
res = abi.hs_filter_result()
h0 = (C.c_double * 4)(); L.hs_debug_host_seconds(e.h, h0)
lat = []
for p, keep in built:
    t = time.perf_counter()
    L.hs_filter(e.h, C.byref(p), None, 0, C.byref(res))
    lat.append(time.perf_counter() - t)
h1 = (C.c_double * 4)(); L.hs_debug_host_seconds(e.h, h1)
n = len(built)
lat.sort()
print('calls', n, 'p50 %.1f us p99 %.1f us mean %.1f us' % (1e6*lat[n//2], 1e6*lat[int(n*0.99)], 1e6*sum(lat)/n))
print('per call: compile+stage %.1f us, device %.1f us, absorb %.1f us, rest(assemble, idle, python) %.1f us' % tuple(
    [1e6*(h1[i]-h0[i])/n for i in range(3)] + [1e6*(sum(lat) - (h1[0]-h0[0]) - (h1[1]-h0[1]) - (h1[2]-h0[2]))/n]))
print('device ms last', e.last_device_ms())
