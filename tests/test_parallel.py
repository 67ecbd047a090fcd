"""Host-side multi-GPU logic on the CPU: sharding and the count/aggregate combine over a world_size-2 gloo group."""
import os
import socket

import pytest

from duckdb_arrow_b200 import parallel as P


@pytest.mark.parametrize("n,world", [(489, 8), (49, 4), (5, 8), (0, 2), (7, 1), (4884, 8)])
def test_batch_range_partitions(n, world):
    ranges = [P.batch_range(n, r, world) for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (a, b), (c, d) in zip(ranges, ranges[1:]):
        assert b == c and a <= b
    sizes = [b - a for a, b in ranges]
    assert max(sizes) - min(sizes) <= 1


def test_file_shard_covers_every_file_once():
    paths = [f"f{i:02d}.arrows" for i in range(64)]
    seen = []
    for r in range(8):
        seen += P.file_shard(paths, r, 8)
    assert sorted(seen) == paths and len(P.file_shard(paths, 3, 8)) == 8
    with pytest.raises(ValueError):
        P.file_shard(paths, 8, 8)


def test_limb_round_trip():
    for v in [0, 1, -1, 2**63, -2**63, 2**127 - 1, -2**127, 123141078228300, -987654321987654321987654321]:
        assert P._from_limbs(P._to_limbs(v)) == v


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = {"rows": 1000 + rank, "selected": 70 * (rank + 1), "sum": (2**100 + 5) if rank == 0 else -(2**99) - 7}
    out = P.combine_counts(mine)
    lo, hi = P.batch_range(489, rank, world)
    out2 = P.combine_counts({"batches": hi - lo})
    q.put((rank, out, out2))
    dist.destroy_process_group()


def test_combine_counts_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, out, out2 in res:
        assert out == {"rows": 2001, "selected": 210, "sum": 2**100 + 5 - 2**99 - 7}
        assert out2 == {"batches": 489}


def test_combine_counts_without_group_is_identity():
    assert P.combine_counts({"rows": 5, "sum": -3}) == {"rows": 5, "sum": -3}


def test_numa_binding_helpers(tmp_path):
    """bind_to_device_numa_node: cpulist parsing, sysfs lookup, and 'unknown node' = no-op (the GPU box is a VM with one
    virtual node and numa_node = -1)."""
    import importlib.util
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("_par", os.path.join(root, "duckdb_arrow_b200", "parallel.py"))
    par = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(par)
    assert par._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert par._parse_cpulist("") == set()
    d = tmp_path / "bus" / "pci" / "devices" / "0000:9c:00.0"
    d.mkdir(parents=True)
    (d / "numa_node").write_text("1\n")
    assert par.device_numa_node("0000:9C:00.0", str(tmp_path)) == 1
    assert par.device_numa_node("0000:aa:00.0", str(tmp_path)) == -1
    info = par.bind_to_device_numa_node(0, str(tmp_path))  # no CUDA device here: best effort, never raises
    assert info["bound"] is False
