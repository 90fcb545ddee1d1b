"""CPU, world_size 2, gloo: the host-side multi-GPU logic (shard by bytes, per-rank compute, variable-length gather,
ordered merge) gives exactly the single-process result. The per-rank compute here is the ORACLE (this is a test of the
sharding plumbing on a GPU-less box); on the GPU box the same function is exercised with the CUDA path
(test_gpu_parity.py::test_sharded_equals_unsharded)."""
import os, socket, sys
import numpy as np, pytest
import torch.multiprocessing as mp
import datasets, orc
from conftest import load_package

def free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p

def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import datasets as ds, orc as orc_
    from conftest import load_package as lp
    sharding = lp().sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = ds.synth_index()
    o = orc_.Index(s["ri"], s["tags0"], 0)
    rs = ds.ragged_reads(s["text"], seed=3, n=400)
    buf, off = orc_.pack_reads(rs)
    def compute(r, of, ml, mo):
        R = o.find_mems(r, of, ml, mo, text=False); R["extensions"] = 0; return R
    out = sharding.find_mems_sharded(compute, buf, off, 12, 1)
    if rank == 0:
        q.put({k: out[k] for k in ("mem_off", "mems", "tag_off", "tags", "n_runs", "n_reads", "n_mems", "n_tags")})
    else:
        assert out is None
    dist.destroy_process_group()

def test_shard_bounds_cover_every_read_once():
    sharding = load_package().sharding
    rng = np.random.default_rng(1)
    for n in (0, 1, 2, 7, 1000):
        lens = rng.integers(0, 300, n); off = np.zeros(n + 1, dtype=np.uint64); off[1:] = np.cumsum(lens)
        for world in (1, 2, 3, 8):
            b = [sharding.shard_bounds(off, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            if n == 1000:
                by = [int(off[h]) - int(off[l]) for l, h in b]
                assert max(by) - min(by) <= 600                       # balanced by bytes up to one read
    off = np.zeros(11, dtype=np.uint64)                               # all-empty reads: balanced by count
    assert [sharding.shard_bounds(off, 2, r) for r in range(2)] == [(0, 5), (5, 10)]

def test_merge_is_identity_on_split_results():
    sharding = load_package().sharding
    s = datasets.synth_index(); o = orc.Index(s["ri"], s["tags0"], 0)
    reads, off = datasets.sample_reads(s["text"], 300, 150, 0.02, seed=5)
    whole = o.find_mems(reads, off, 20, 1, text=False)
    parts = []
    for r in range(3):
        rr, oo, lo, hi = sharding.shard_reads(reads, off, 3, r)
        parts.append(o.find_mems(rr, oo, 20, 1, text=False))
    m = sharding.merge_results(parts)
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(m[k], whole[k]), k
    assert (m["n_reads"], m["n_mems"], m["n_tags"]) == (whole["n_reads"], whole["n_mems"], whole["n_tags"])

def test_world_size_2_gloo_equals_single_process():
    s = datasets.synth_index(); o = orc.Index(s["ri"], s["tags0"], 0)
    rs = datasets.ragged_reads(s["text"], seed=3, n=400); buf, off = orc.pack_reads(rs)
    whole = o.find_mems(buf, off, 12, 1, text=False)
    ctx = mp.get_context("spawn"); q = ctx.Queue(); port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = q.get(timeout=240)
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for k in ("mem_off", "mems", "tag_off", "tags", "n_runs"): assert np.array_equal(got[k], whole[k]), k
    assert got["n_reads"] == whole["n_reads"] == len(rs) and got["n_mems"] == whole["n_mems"] > 0


def _bcast_worker(rank, world, port, q):
    import torch, torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from conftest import load_package as lp
    pkg = lp(); sharding = pkg.sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    runs = None
    if rank == 0:
        d = pkg.pangraph.build(4000, 5, seed=2, device="cpu", snp=1 / 100, indel=1 / 300, n_stretch=2); runs = d["dev_runs"]
    got, nbytes = sharding.broadcast_runs(runs, torch.device("cpu"))
    q.put((rank, {k: v.numpy().copy() for k, v in got.items()}, nbytes))
    dist.barrier(); dist.destroy_process_group()

def test_world_size_2_gloo_broadcast_of_runs():
    """bench.py's "index built once per node": rank 0 derives the run-length BWT + tag runs, every rank receives them."""
    pkg = load_package()
    want = pkg.pangraph.build(4000, 5, seed=2, device="cpu", snp=1 / 100, indel=1 / 300, n_stretch=2)["dev_runs"]
    ctx = mp.get_context("spawn"); q = ctx.Queue(); port = free_port()
    procs = [ctx.Process(target=_bcast_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = dict((r, (a, b)) for r, a, b in (q.get(timeout=240) for _ in range(2)))
    [p.join(timeout=120) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for r in (0, 1):
        for k in ("sym", "len", "tag_val", "tag_len"): assert np.array_equal(got[r][0][k], want[k].numpy()), (r, k)
        assert got[r][1] == sum(v.numel() * v.element_size() for v in want.values())
