"""Diagnostic (not part of the product): host<->device copy bandwidth per rank when N ranks copy at once, for torch
pinned memory, lsg_host_alloc default and write-combined memory, with and without CPU affinity to the GPU's local CPUs.
Run under torchrun; each rank appends a line to gpurun_out/pcie_probe.txt."""
import ctypes as C
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_slo_ebpf_toolkit_b200._lib import HostBuffer  # noqa: E402

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
props = torch.cuda.get_device_properties(local)
bdf = "%04x:%02x:%02x.0" % (getattr(props, "pci_domain_id", 0), getattr(props, "pci_bus_id", 0), getattr(props, "pci_device_id", 0))
sysd = "/sys/bus/pci/devices/" + bdf
info = {}
for f in ("numa_node", "local_cpulist", "current_link_speed", "current_link_width"):
    try:
        info[f] = open(os.path.join(sysd, f)).read().strip()
    except OSError as e:
        info[f] = "?"
UP, DOWN = 352 << 20, 135 << 20
dev_up = torch.empty(UP, dtype=torch.uint8, device="cuda")
dev_dn = torch.empty(DOWN, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
rt = C.CDLL("libcudart.so.12")
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def measure(up_ptr, dn_ptr, reps=8):
    out = {}
    for mode in ("up", "down", "both"):
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            if mode in ("up", "both"):
                rt.cudaMemcpyAsync(dev_up.data_ptr(), up_ptr, UP, 1, s_up.cuda_stream)
            if mode in ("down", "both"):
                rt.cudaMemcpyAsync(dn_ptr, dev_dn.data_ptr(), DOWN, 2, s_dn.cuda_stream)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        nbytes = (UP if mode in ("up", "both") else 0) + (DOWN if mode in ("down", "both") else 0)
        out[mode] = round(nbytes * reps / dt / 1e9, 1)
        out[mode + "_ms"] = round(dt / reps * 1e3, 2)
    return out


res = {}
tp_up = torch.empty(UP, dtype=torch.uint8, pin_memory=True)
tp_dn = torch.empty(DOWN, dtype=torch.uint8, pin_memory=True)
tp_up.fill_(1)
res["torch_pinned"] = measure(tp_up.data_ptr(), tp_dn.data_ptr())
hb_up, hb_dn = HostBuffer(UP), HostBuffer(DOWN)
hb_up.array[:] = 1
res["lsg_default"] = measure(hb_up.ptr, hb_dn.ptr)
wc_up = HostBuffer(UP, write_combined=True)
wc_up.array[:] = 1
res["lsg_wc_up"] = measure(wc_up.ptr, hb_dn.ptr)
# affinity to the GPU's local CPUs, then fresh allocations (first touch on the local node)
aff = "unchanged"
try:
    cl = info["local_cpulist"]
    cpus = set()
    for part in cl.split(","):
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    allowed = os.sched_getaffinity(0)
    if cpus and cpus & allowed and (cpus & allowed) != allowed:
        os.sched_setaffinity(0, cpus & allowed)
        aff = "%d cpus" % len(cpus & allowed)
except Exception as e:  # noqa: BLE001
    aff = "error %s" % e
af_up, af_dn = HostBuffer(UP), HostBuffer(DOWN)
af_up.array[:] = 1
af_dn.array[:] = 0
res["affinity(%s)" % aff] = measure(af_up.ptr, af_dn.ptr)
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/pcie_probe.txt", "a") as fh:
    fh.write("rank %d/%d gpu %s %s ncpu=%d %s\n" % (rank, world, bdf, info, len(os.sched_getaffinity(0)), res))
barrier()
if world > 1:
    dist.destroy_process_group()
