"""Multi-GPU plumbing: one process per GPU (torch.distributed), record batches / files sharded across ranks, and
the ONE collective step of the path — combining per-GPU row counts and aggregate partials.

Record batches are self-contained (dictionaries are replicated per file), so the data path needs no exchange:
rank r decodes the contiguous batch range batch_range(n, r, world) of a file, or the files file_shard(paths, r,
world) of a glob (the reference scans one thread per file: src/file_scanner/arrow_multi_file_info.cpp:77-86).
Only {rows, selected rows, aggregate partials} cross GPUs: an all_reduce(SUM) of a few int64 values (NCCL over
NVLink on the GPU box, gloo in the CPU tests)."""
from __future__ import annotations

LIMB_BITS = 32
N_LIMBS = 4  # 128-bit two's complement (DuckDB HUGEINT aggregates)


def batch_range(n_batches: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced partition: output order = rank-major = file order (matches a single-file scan)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    return n_batches * rank // world, n_batches * (rank + 1) // world


def file_shard(paths: list[str], rank: int, world: int) -> list[str]:
    """Whole files of a (sorted) glob, round-robin so that sizes balance when files are alike."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    return list(paths[rank::world])


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def device_numa_node(pci_bus_id: str, sysfs: str = "/sys") -> int:
    """NUMA node the PCIe device hangs off (-1: unknown / single node / virtualised)."""
    try:
        with open(f"{sysfs}/bus/pci/devices/{pci_bus_id.lower()}/numa_node") as f:
            return int(f.read().strip())
    except (OSError, ValueError):
        return -1


def bind_to_device_numa_node(device: int, sysfs: str = "/sys") -> dict:
    """One process per GPU: run this rank's host threads on the cores of the NUMA node its GPU hangs off, BEFORE the
    pinned buffers are allocated, so that first-touch places them in that node's memory and the rank's H2D / D2H
    traffic stays off the inter-socket link.  The end-to-end leg streams ~8.7 GB per step and GPU through host memory;
    nothing else of the path is shared between ranks.  Best effort: returns what it did, never raises."""
    import os
    info = {"device": device, "numa_node": -1, "bound": False}
    try:
        import torch
        p = torch.cuda.get_device_properties(device)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        info["pci"] = bdf
        node = device_numa_node(bdf, sysfs)
        info["numa_node"] = node
        if node < 0:
            return info
        with open(f"{sysfs}/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read()) & set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["bound"] = True
            info["cpus"] = len(cpus)
    except Exception as e:  # noqa: BLE001 - topology files differ between kernels / containers
        info["error"] = repr(e)
    return info


def _to_limbs(v: int) -> list[int]:
    u = v & ((1 << (LIMB_BITS * N_LIMBS)) - 1)
    return [(u >> (LIMB_BITS * i)) & ((1 << LIMB_BITS) - 1) for i in range(N_LIMBS)]


def _from_limbs(limbs: list[int]) -> int:
    u = 0
    for i, l in enumerate(limbs):
        u += int(l) << (LIMB_BITS * i)
    u &= (1 << (LIMB_BITS * N_LIMBS)) - 1
    return u - (1 << (LIMB_BITS * N_LIMBS)) if u >> (LIMB_BITS * N_LIMBS - 1) else u


def combine_counts(values: dict[str, int], device=None, group=None) -> dict[str, int]:
    """all_reduce(SUM) of named integers (each may be a full 128-bit aggregate).  Each value travels as four 32-bit
    limbs in int64 lanes, so the limb sums cannot overflow for any realistic world size; carries are resolved on
    the host.  Without an initialised process group (single GPU) the values are returned unchanged."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return dict(values)
    keys = sorted(values)
    flat = []
    for k in keys:
        flat += _to_limbs(int(values[k]))
    t = torch.tensor(flat, dtype=torch.int64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    out = t.cpu().tolist()
    return {k: _from_limbs(out[i * N_LIMBS:(i + 1) * N_LIMBS]) for i, k in enumerate(keys)}


def sharded_scan(reader, rank: int, world: int, device: int = 0, **kw):
    """The rank's share of a single-file scan: ArrowScan over batch_range(reader.num_batches, rank, world)."""
    from .api import ArrowScan
    lo, hi = batch_range(reader.num_batches, rank, world)
    return ArrowScan(reader, device=device, batch_lo=lo, batch_hi=hi, **kw)


def sharded_filter_aggregate(reader, predicates, col_a: str, col_b: str, rank: int, world: int, device: int = 0):
    """BASELINE config 3: pushed-down filters + the Q6-style sum(a*b), batches sharded across GPUs.
    Returns the GLOBAL {rows, selected, sum} after the one collective step."""
    scan = sharded_scan(reader, rank, world, device)
    rows = sum(reader.batch_info(b).rows for b in range(scan.batch_lo, scan.batch_hi))
    selected, total = 0, 0
    if scan.batch_hi > scan.batch_lo:
        scan.upload()
        scan.decode()
        sel, count, tot = scan.filter(predicates)
        total = scan.sum_product(col_a, col_b, sel, count)
        scan.finish()
        selected = int(tot.item())
    return combine_counts({"rows": rows, "selected": selected, "sum": total}, device=f"cuda:{device}")
