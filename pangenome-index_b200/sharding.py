"""One process per GPU: shard a read batch across ranks, run the hot path on each rank's replica of the index, and
merge the per-rank results back into read order. Reads are independent, so there is NO collective
on the data path; `torch.distributed` is used only to hand the per-rank results to rank 0 (variable-length gather).

The reference has no counterpart (its loop is one thread, src/find_mems.cpp:94-139); the C++ driver
host/find_mems.cpp does the same thing with threads inside one process.
"""
import numpy as np

RESULT_ARRAYS = ("mem_off", "mems", "tag_off", "tags", "n_runs")

def shard_bounds(read_off, world, rank):
    """Contiguous range [lo, hi) of reads for `rank`, balanced by BYTES (ragged reads: work follows bases, not reads).
    Every read belongs to exactly one rank; ranges are ordered by rank so concatenation restores read order."""
    read_off = np.asarray(read_off, dtype=np.uint64)
    n = len(read_off) - 1
    if n <= 0:
        return 0, 0
    total = int(read_off[-1])
    if total == 0:                                   # all reads empty: balance by count
        return (n * rank) // world, (n * (rank + 1)) // world
    cuts = [int(np.searchsorted(read_off, np.uint64((total * r) // world), side="left")) for r in range(world + 1)]
    cuts[0] = 0; cuts[-1] = n
    for i in range(1, world + 1):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return min(cuts[rank], n), min(cuts[rank + 1], n)

def shard_reads(reads, read_off, world, rank):
    """-> (reads slice, offsets rebased to 0, lo, hi)."""
    reads = np.asarray(reads, dtype=np.uint8); read_off = np.asarray(read_off, dtype=np.uint64)
    lo, hi = shard_bounds(read_off, world, rank)
    b0, b1 = int(read_off[lo]), int(read_off[hi])
    return reads[b0:b1], (read_off[lo:hi + 1] - read_off[lo]).astype(np.uint64), lo, hi

def merge_results(parts):
    """Concatenate per-shard results (dicts as returned by capi.Index.find_mems, in shard order) into one result:
    offsets are rebased, payload arrays are concatenated, counters are summed."""
    out = dict(n_reads=0, n_mems=0, n_tags=0, extensions=0)
    mem_off = [np.zeros(1, dtype=np.uint64)]; tag_off = [np.zeros(1, dtype=np.uint64)]; mems = []; tags = []; n_runs = []
    for p in parts:
        mem_off.append(np.asarray(p["mem_off"][1:], dtype=np.uint64) + np.uint64(out["n_mems"]))
        tag_off.append(np.asarray(p["tag_off"][1:], dtype=np.uint64) + np.uint64(out["n_tags"]))
        mems.append(np.asarray(p["mems"], dtype=np.uint64).reshape(-1, 4)); tags.append(np.asarray(p["tags"], dtype=np.uint64))
        n_runs.append(np.asarray(p["n_runs"], dtype=np.uint64))
        for k in ("n_reads", "n_mems", "n_tags"): out[k] += int(p[k])
        out["extensions"] += int(p.get("extensions", 0))
    out["mem_off"] = np.concatenate(mem_off); out["tag_off"] = np.concatenate(tag_off)
    out["mems"] = np.concatenate(mems) if mems else np.zeros((0, 4), dtype=np.uint64)
    out["tags"] = np.concatenate(tags) if tags else np.zeros(0, dtype=np.uint64)
    out["n_runs"] = np.concatenate(n_runs) if n_runs else np.zeros(0, dtype=np.uint64)
    return out

def find_mems_sharded(compute, reads, read_off, min_len, min_occ, group=None, dst=0):
    """Every rank calls this with the SAME batch. `compute(reads, off, min_len, min_occ) -> result dict` runs on the
    rank's own GPU (capi.Index(...).find_mems bound to that rank's replica). Returns the merged result on rank `dst`,
    None elsewhere. Without an initialised process group it degenerates to a single shard."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return merge_results([compute(reads, read_off, min_len, min_occ)])
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    r, o, lo, hi = shard_reads(reads, read_off, world, rank)
    local = compute(r, o, min_len, min_occ)
    local = {k: local[k] for k in RESULT_ARRAYS + ("n_reads", "n_mems", "n_tags", "extensions") if k in local}
    gathered = [None] * world if rank == dst else None
    dist.gather_object(local, gathered, dst=dst, group=group)
    return merge_results(gathered) if rank == dst else None


RUN_ARRAYS = (("sym", "uint8"), ("len", "int64"), ("tag_val", "int64"), ("tag_len", "int64"))

def broadcast_runs(runs, device, src=0, group=None):
    """The index is built ONCE per node: rank `src` holds the run-length BWT (`sym`, `len`) and the run-length tag array
    (`tag_val`, `tag_len`) as tensors on its device, every other rank passes None; all ranks return the four tensors on
    `device` (NCCL over NVLink on the GPU box, gloo in the CPU tests). Each rank then constructs its own replica from them
    (pgi_index_create_from_runs). -> (runs dict, bytes received per rank)"""
    import torch, torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return runs, 0
    rank = dist.get_rank(group)
    meta = torch.zeros(2, dtype=torch.int64, device=device)
    if rank == src: meta[0] = runs["sym"].numel(); meta[1] = runs["tag_val"].numel()
    dist.broadcast(meta, src, group=group)
    n_runs, n_tag = int(meta[0]), int(meta[1])
    if rank != src:
        runs = {k: torch.empty(n_runs if k in ("sym", "len") else n_tag, dtype=getattr(torch, dt), device=device) for k, dt in RUN_ARRAYS}
    for k, _ in RUN_ARRAYS: dist.broadcast(runs[k], src, group=group)
    return runs, sum(runs[k].numel() * runs[k].element_size() for k, _ in RUN_ARRAYS)
