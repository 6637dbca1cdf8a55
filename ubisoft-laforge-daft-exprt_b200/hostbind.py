"""CPU / NUMA placement of a rank next to its GPU (host-side runtime of the bulk path).

One process drives one GPU.  Its pinned staging blocks are first-touched by, and its collate / writer threads run on,
host cores; when eight ranks share one box, leaving all of them on the default affinity mask means every rank's staging
memory can land on one NUMA node and every rank's helper threads migrate over all cores (round-1 8-GPU record: all GPUs
reported CPU affinity 0-31 / NUMA 0, host-buffer scaling efficiency 0.64).  ``bind_to_gpu`` pins the calling process:

* to the cores sysfs lists as local to the GPU's PCI device (``/sys/bus/pci/devices/<bdf>/local_cpulist``) when that is a
  strict subset of the cores the process may use (a real multi-socket topology), then
* inside that set — or inside the whole allowed set when the box exposes one NUMA node — to the ``local_rank``-th of
  ``local_world`` equal contiguous slices, so that ranks never share cores.

It must run before the first pinned allocation (first touch decides the page's node) and returns a record for logs.
Linux only; every failure is reported in the record and leaves the process unbound.
"""
import os


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_pci_bdf(device_index):
    import torch
    p = torch.cuda.get_device_properties(device_index)
    return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"


def _sysfs(bdf, name):
    try:
        with open(f"/sys/bus/pci/devices/{bdf}/{name}") as f:
            return f.read().strip()
    except OSError:
        return None


def slice_for_rank(cpus, local_rank, local_world):
    """The ``local_rank``-th of ``local_world`` contiguous, near-equal slices of the sorted core list (never empty)."""
    cpus = sorted(cpus)
    if local_world <= 1 or len(cpus) < local_world:
        return cpus
    per, extra = divmod(len(cpus), local_world)
    start = local_rank * per + min(local_rank, extra)
    return cpus[start:start + per + (1 if local_rank < extra else 0)]


def bind_to_gpu(device_index, local_rank=0, local_world=1, set_torch_threads=True):
    rec = {"bound": False, "gpu": int(device_index), "local_rank": int(local_rank), "local_world": int(local_world)}
    try:
        allowed = set(os.sched_getaffinity(0))
        rec["allowed_cpus"] = len(allowed)
        bdf = gpu_pci_bdf(device_index)
        rec["pci"] = bdf
        node = _sysfs(bdf, "numa_node")
        local = _sysfs(bdf, "local_cpulist")
        rec["numa_node"] = int(node) if node not in (None, "") else None
        pool = allowed
        if local:
            near = _parse_cpulist(local) & allowed
            if near and near != allowed:
                pool = near
                rec["numa_local"] = True
        # ranks whose GPUs share a pool split it; with one pool for all (single node / VM) every rank gets its slice
        mine = slice_for_rank(pool, local_rank, local_world) if pool == allowed else sorted(pool)
        if pool != allowed:
            # how many ranks share this GPU's node is not visible from here: split the node by the ranks that can be on it
            per_node = max(1, local_world * len(pool) // max(1, len(allowed)))
            mine = slice_for_rank(pool, local_rank % per_node, per_node)
        os.sched_setaffinity(0, mine)
        rec.update(bound=True, cpus=f"{mine[0]}-{mine[-1]}" if mine == list(range(mine[0], mine[-1] + 1)) else mine, n_cpus=len(mine))
        if set_torch_threads:
            import torch
            torch.set_num_threads(max(1, len(mine)))
    except Exception as e:          # noqa: BLE001 - placement is an optimisation, never a failure
        rec["error"] = f"{type(e).__name__}: {e}"
    return rec
