import os, sys, glob, time
import torch
def gpu_numa():
    p = torch.cuda.get_device_properties(0)
    bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    try:
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
    except Exception as ex:
        node = None
    return bdf, node
def cpus_of(node):
    try:
        s = open(f"/sys/devices/system/node/node{node}/cpulist").read().strip()
    except Exception:
        return None
    out = []
    for part in s.split(","):
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out
def bw(sizes, reps=200, graph=False):
    src = [torch.empty(s, dtype=torch.uint8).pin_memory() for s in sizes]
    for t in src: t.fill_(1)
    dst = [torch.empty(s, dtype=torch.uint8, device="cuda") for s in sizes]
    st = torch.cuda.Stream()
    def run():
        for d, s in zip(dst, src): d.copy_(s, non_blocking=True)
    with torch.cuda.stream(st):
        for _ in range(5): run()
        torch.cuda.synchronize()
        if graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                for _ in range(24): run()
            fn, k = g.replay, 24
        else:
            fn, k = run, 1
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        fn(); torch.cuda.synchronize()
        e0.record(st)
        for _ in range(reps // k + 1): fn()
        e1.record(st)
        e1.synchronize()
        n = (reps // k + 1) * k
    return sum(sizes) * n / (e0.elapsed_time(e1) * 1e-3) / 1e9
print("nodes:", [os.path.basename(p) for p in glob.glob("/sys/devices/system/node/node*")], "affinity:", len(os.sched_getaffinity(0)), "cpus")
bdf, node = gpu_numa()
print("gpu", bdf, "numa", node)
parts = [212992, 1572864, 835584]
for tag in ("default", "local"):
    if tag == "local":
        c = cpus_of(node) if node is not None and node >= 0 else None
        if not c:
            print("no numa info; skipping local"); break
        os.sched_setaffinity(0, set(c) & os.sched_getaffinity(0) or set(c))
        print("bound to", len(os.sched_getaffinity(0)), "cpus of node", node)
    for g in (False, True):
        print(tag, "graph" if g else "eager", "split3 %.1f GB/s" % bw(parts, graph=g), "single %.1f GB/s" % bw([sum(parts)], graph=g), "big64MB %.1f GB/s" % bw([64 << 20], reps=20, graph=False))
