"""Multi-GPU sharding of the `stoat vcf` hot path: one process per GPU.

The reference has no multi-process mode; its unit of parallelism is the snarl inside a chromosome chunk
(`#pragma omp parallel for`, src/snarl_analyzer.cpp:165-172) and nothing is shared between snarls except the
read-only matrix and the phenotype.  Two shardings follow from that:

snarl-sharded (default)
    Contiguous snarl ranges per rank, balanced by checked edges.  A rank receives only the matrix rows its
    snarls reference (row indices are remapped to a compact local matrix) and a replica of the phenotype.
    There is NO data-path collective: each rank runs the ordinary batch call and the host concatenates the
    per-rank result arrays in rank order (`merge_snarl_results`).

sample-sharded
    Every rank holds all rows but only a range of haplotype columns (cut at multiples of 128 columns, so both
    haplotypes of a sample and every 16-byte chunk stay on one rank) plus its slice of the phenotype.  Counts
    are sums over haplotypes, so ONE exchange is needed: all-reduce(sum) of the uint32 [2P] count buffer
    between the count pass (K1) and the statistics pass (K2).  On GPUs that is an NCCL all-reduce over NVLink
    applied in place to the context's device buffer; on CPU (tests, gloo) the same driver runs with host tensors.

Everything here is host plumbing (numpy + torch.distributed); the statistics come from the CUDA library.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .api import SnarlPaths

MISSING_EDGE = 0xFFFFFFFF


# ---------------------------------------------------------------------------------------------------
# snarl-sharded
# ---------------------------------------------------------------------------------------------------
def plan_snarl_shards(paths: SnarlPaths, world: int) -> list[tuple[int, int]]:
    """Contiguous snarl ranges [s0, s1) per rank, balanced by work = checked edges + paths (the matrix rows a
    snarl streams are proportional to its checked edges; a path without edges still costs a table column)."""
    S = paths.n_snarls
    if world <= 0:
        raise ValueError("world must be positive")
    spo = paths.snarl_path_off.astype(np.int64)
    peo = paths.path_edge_off.astype(np.int64)
    # cumulative work at snarl boundaries
    work = peo[spo] + spo
    total = int(work[-1]) if S else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        s = int(np.searchsorted(work, target, side="left"))
        cuts.append(min(max(s, cuts[-1]), S))
    cuts.append(S)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


@dataclass
class SnarlShard:
    s0: int
    s1: int
    paths: SnarlPaths        # local CSR, edge_row remapped to local rows
    rows: np.ndarray         # int64 [n_local_rows]: global row of each local row (ascending)
    bits: np.ndarray | None  # local matrix in the reference layout (None until `attach_rows`)
    n_rows: int
    n_cols: int


def slice_paths(paths: SnarlPaths, s0: int, s1: int) -> SnarlPaths:
    p0, p1 = int(paths.snarl_path_off[s0]), int(paths.snarl_path_off[s1])
    e0, e1 = int(paths.path_edge_off[p0]), int(paths.path_edge_off[p1])
    return SnarlPaths(paths.snarl_path_off[s0:s1 + 1] - np.uint64(p0),
                      paths.path_edge_off[p0:p1 + 1] - np.uint64(e0), paths.edge_row[e0:e1].copy())


def gather_rows(bits: np.ndarray, n_cols: int, rows: np.ndarray) -> np.ndarray:
    """Rows `rows` of a reference-layout bit matrix (bit = row * n_cols + col, LSB first, src/matrix.cpp:67-84)
    as a new reference-layout matrix with len(rows) rows."""
    rows = np.asarray(rows, dtype=np.int64)
    if rows.size == 0:
        return np.zeros(0, np.uint8)
    if n_cols % 8 == 0:  # rows are byte aligned
        rb = n_cols // 8
        return np.ascontiguousarray(bits[: (int(rows.max()) + 1) * rb].reshape(-1, rb)[rows]).reshape(-1)
    out = np.zeros((rows.size, n_cols), dtype=np.uint8)
    for k, r in enumerate(rows):  # unaligned rows: bit-level extraction, row by row
        b0 = int(r) * n_cols
        seg = np.unpackbits(bits[b0 // 8:(b0 + n_cols + 7) // 8 + 1], bitorder="little")
        out[k] = seg[b0 % 8:b0 % 8 + n_cols]
    return np.packbits(out.reshape(-1), bitorder="little")


def extract_snarl_shard(bits: np.ndarray | None, n_rows: int, n_cols: int, paths: SnarlPaths, s0: int,
                        s1: int) -> SnarlShard:
    """The self-contained problem of snarls [s0, s1): local CSR + the compact matrix of the rows they use.
    Rows shared with a neighbouring shard are duplicated (they are read-only)."""
    local = slice_paths(paths, s0, s1)
    er = local.edge_row.astype(np.int64)
    present = er != MISSING_EDGE
    if np.any(er[present] >= n_rows):
        raise ValueError("edge_row holds a row index >= n_rows")
    rows = np.unique(er[present])
    remap = er.copy()
    remap[present] = np.searchsorted(rows, er[present])
    local = SnarlPaths(local.snarl_path_off, local.path_edge_off, remap.astype(np.uint32))
    local_bits = gather_rows(bits, n_cols, rows) if bits is not None else None
    return SnarlShard(s0, s1, local, rows, local_bits, int(rows.size), n_cols)


PER_PATH_KEYS = ("g0", "g1", "allele_paths")


def merge_snarl_results(parts: list[dict]) -> dict:
    """Concatenate per-rank result dicts (rank order = snarl order).  Per-path and per-snarl arrays are both
    plain concatenations because shards are contiguous snarl ranges."""
    if not parts:
        return {}
    return {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}


def run_snarl_sharded(ctx, trait: str, bits, n_rows: int, n_cols: int, paths: SnarlPaths, phenotype, rank: int,
                      world: int, group=None, thresholds=(3, 5, 0.05), covariates=None) -> dict | None:
    """One rank's part of a snarl-sharded run on its own GPU context; rank 0 returns the merged result.
    No data-path collective: the only communication is the host gather of the result arrays."""
    import torch.distributed as dist

    s0, s1 = plan_snarl_shards(paths, world)[rank]
    sh = extract_snarl_shard(bits, n_rows, n_cols, paths, s0, s1)
    ctx.set_matrix(sh.bits if sh.n_rows else np.zeros(1, np.uint8), sh.n_rows, n_cols)
    mi, mh, maf = thresholds
    if trait == "binary":
        ctx.set_binary_phenotype(phenotype)
        part = ctx.binary_batch(sh.paths, mi, mh, maf)
    else:
        ctx.set_quantitative_phenotype(phenotype)
        ctx.set_covariates(covariates)
        part = ctx.quantitative_batch(sh.paths, mi, mh, maf)
    if world == 1:
        return part
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(part, gathered, dst=0, group=group)
    return merge_snarl_results(gathered) if rank == 0 else None


# ---------------------------------------------------------------------------------------------------
# sample-sharded
# ---------------------------------------------------------------------------------------------------
def plan_column_shards(n_cols: int, world: int, align: int = 128) -> list[tuple[int, int]]:
    """Haplotype column ranges [c0, c1) per rank, cut at multiples of `align` columns (128 = one 16-byte chunk of
    a repacked row; even, so both haplotypes of a sample stay together).  Trailing ranks may be empty."""
    if n_cols % 2:
        raise ValueError("n_cols must be even")
    chunks = (n_cols + align - 1) // align
    cuts = [min(n_cols, align * ((chunks * r) // world)) for r in range(world)] + [n_cols]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def split_columns(bits: np.ndarray, n_rows: int, n_cols: int, c0: int, c1: int) -> np.ndarray:
    """Columns [c0, c1) of every row as a reference-layout matrix with c1 - c0 columns."""
    w = c1 - c0
    if w <= 0 or n_rows == 0:
        return np.zeros(0, np.uint8)
    if n_cols % 8 == 0 and c0 % 8 == 0 and w % 8 == 0:
        rb = n_cols // 8
        return np.ascontiguousarray(bits[: n_rows * rb].reshape(n_rows, rb)[:, c0 // 8:c1 // 8]).reshape(-1)
    dense = np.unpackbits(bits[: (n_rows * n_cols + 7) // 8], bitorder="little")[: n_rows * n_cols]
    return np.packbits(dense.reshape(n_rows, n_cols)[:, c0:c1].reshape(-1), bitorder="little")


class _DevU32:
    """__cuda_array_interface__ view of a device uint32 buffer as int32 (two's-complement sums are identical),
    so that torch / NCCL can all-reduce the context's count buffer in place."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 3}


def allreduce_counts_inplace(ctx, group=None):
    """NCCL all-reduce(sum) applied to the context's device count buffer [g0 | g1] (uint32 [2P])."""
    import torch
    import torch.distributed as dist

    ptr, n = ctx.binary_counts_dev()
    if n == 0:
        return
    t = torch.as_tensor(_DevU32(ptr, n), device=torch.device("cuda", ctx.device))
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    torch.cuda.synchronize(ctx.device)


def run_binary_sample_sharded(ctx, bits, n_rows: int, n_cols: int, paths: SnarlPaths, is_case, rank: int, world: int,
                              group=None, thresholds=(3, 5, 0.05)) -> dict:
    """One rank's part of a sample-sharded binary run: count pass on the local columns, ONE all-reduce of the
    counts, statistics on the summed counts (replicated: every rank returns the full result).

    A path without checked edges matches every haplotype (src/snarl_analyzer.cpp:343-354): each rank counts its
    own columns for it, so the sum is still the global count.  A rank whose column range is empty contributes
    zeros and takes part in the collective."""
    c0, c1 = plan_column_shards(n_cols, world)[rank]
    mi, mh, maf = thresholds
    if c1 > c0:
        local = split_columns(bits, n_rows, n_cols, c0, c1)
        ctx.set_matrix(local, n_rows, c1 - c0)
        ctx.set_binary_phenotype(np.asarray(is_case, np.uint8)[c0 // 2:c1 // 2])
        ctx.stage_paths(paths)
        ctx.binary_counts_staged()
    else:
        raise ValueError("sample-sharded run needs at least 128 columns per rank")
    if world > 1:
        allreduce_counts_inplace(ctx, group)
    ctx.binary_stats_staged(mi, mh, maf)
    return ctx.binary_fetch()


# ---------------------------------------------------------------------------------------------------
# sample-sharded, quantitative
# ---------------------------------------------------------------------------------------------------
P_DEFERRED = 7  # STOAT_P_DEFERRED (include/stoat_gpu.h)


def phenotype_summary(y) -> tuple[float, float, float, float]:
    """mean, sum (y - mean), sum (y - mean)^2, max |y - mean| of the WHOLE phenotype: what every rank of a
    sample-sharded run passes to set_quantitative_phenotype_shard (same numbers on every rank: computed here in
    extended precision exactly as the library does for an unsharded phenotype)."""
    a = np.asarray(y, dtype=np.float64)
    if a.size == 0:
        return 0.0, 0.0, 0.0, 0.0
    mean = float(np.sum(a.astype(np.longdouble)) / np.longdouble(a.size))
    v = a - mean
    vl = v.astype(np.longdouble)
    return mean, float(np.sum(vl)), float(np.sum(vl * vl)), float(np.max(np.abs(v)))


class _DevF64:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3}


def allreduce_partials_inplace(ctx, group=None):
    """The ONE exchange of a sample-sharded quantitative run: NCCL all-reduce(sum) of the per-snarl class counts
    (uint32 [7 S] per haplotype slice) and sums (float64 [3 S]) in place on the context's device buffers."""
    import torch
    import torch.distributed as dist

    pi, ni, pd, nd = ctx.quantitative_partials_dev()
    dev = torch.device("cuda", ctx.device)
    if ni:
        dist.all_reduce(torch.as_tensor(_DevU32(pi, ni), device=dev), op=dist.ReduceOp.SUM, group=group)
    if nd:
        dist.all_reduce(torch.as_tensor(_DevF64(pd, nd), device=dev), op=dist.ReduceOp.SUM, group=group)
    torch.cuda.synchronize(ctx.device)


def deferred_snarls(result: dict) -> np.ndarray:
    """snarls a sample-sharded solve could not finish from the summed two-path partials"""
    return np.flatnonzero(np.asarray(result["code"]) == P_DEFERRED)


def gather_deferred_rows(local_bits: np.ndarray, n_rows: int, local_cols: int, paths: SnarlPaths, todo: np.ndarray,
                         col_ranges: list[tuple[int, int]], allgather) -> tuple[np.ndarray, int, SnarlPaths]:
    """Whole rows of the snarls in `todo`, rebuilt from every rank's column slice.

    `allgather(obj) -> list` is the collective (dist.all_gather_object in a real run).  Returns a compact
    reference-layout matrix (rows x all columns), its row count and the sub-batch whose edge_row indices point into it."""
    sub = select_snarls(paths, todo)
    er = sub.edge_row
    rows = np.unique(er[er != MISSING_EDGE])
    mine = gather_rows(local_bits, local_cols, rows) if local_cols else np.zeros(0, np.uint8)
    parts = allgather((mine, local_cols))
    n_cols = sum(c1 - c0 for c0, c1 in col_ranges)
    dense = np.zeros((rows.size, n_cols), dtype=bool)
    for (c0, c1), (chunk, w) in zip(col_ranges, parts):
        if c1 <= c0:
            continue
        assert w == c1 - c0
        d = np.unpackbits(np.asarray(chunk, np.uint8), bitorder="little")[: rows.size * w].reshape(rows.size, w)
        dense[:, c0:c1] = d.astype(bool)
    remap = np.full(int(rows.max()) + 1 if rows.size else 1, MISSING_EDGE, dtype=np.uint32)
    remap[rows] = np.arange(rows.size, dtype=np.uint32)
    er2 = er.copy()
    ok = er != MISSING_EDGE
    er2[ok] = remap[er[ok]]
    bits = np.packbits(dense.reshape(-1), bitorder="little")
    return bits, int(rows.size), SnarlPaths(sub.snarl_path_off, sub.path_edge_off, er2)


def _ranges(starts: np.ndarray, lengths: np.ndarray) -> np.ndarray:
    """concatenation of arange(starts[i], starts[i] + lengths[i]) without a Python loop"""
    total = int(lengths.sum())
    if total == 0:
        return np.zeros(0, np.int64)
    ends = np.cumsum(lengths)
    out = np.ones(total, np.int64)
    first = ends - lengths
    nz = lengths > 0
    out[first[nz]] = starts[nz] - np.concatenate([[0], (starts + lengths)[nz][:-1] - 1])
    return np.cumsum(out)


def select_snarls(paths: SnarlPaths, idx: np.ndarray) -> SnarlPaths:
    """The CSR of the snarls `idx` (in that order); vectorised (1M snarls x a few per cent deferred is 70k snarls)."""
    spo = np.asarray(paths.snarl_path_off, dtype=np.int64)
    peo = np.asarray(paths.path_edge_off, dtype=np.int64)
    er = np.asarray(paths.edge_row, dtype=np.uint32)
    idx = np.asarray(idx, dtype=np.int64)
    n_paths = spo[idx + 1] - spo[idx]
    new_spo = np.concatenate([[0], np.cumsum(n_paths)]).astype(np.uint64)
    path_ids = _ranges(spo[idx], n_paths)
    n_edges = peo[path_ids + 1] - peo[path_ids]
    new_peo = np.concatenate([[0], np.cumsum(n_edges)]).astype(np.uint64)
    edges = er[_ranges(peo[path_ids], n_edges)] if path_ids.size else np.zeros(0, np.uint32)
    return SnarlPaths(new_spo, new_peo, edges.astype(np.uint32))


def merge_deferred(result: dict, todo: np.ndarray, replay: dict, paths: SnarlPaths) -> dict:
    """Write the replayed snarls' rows over the deferred ones (per-snarl arrays by snarl, allele_paths by path)."""
    spo = np.asarray(paths.snarl_path_off, dtype=np.int64)
    todo = np.asarray(todo, dtype=np.int64)
    out = {k: np.array(v, copy=True) for k, v in result.items()}
    for k, v in replay.items():
        if k not in out:
            continue
        if k == "allele_paths":
            out[k][_ranges(spo[todo], spo[todo + 1] - spo[todo])] = v
        else:
            out[k][todo] = v
    return out


def remap_rows(sub: SnarlPaths) -> tuple[np.ndarray, SnarlPaths]:
    """(distinct matrix rows of `sub`, `sub` with edge_row pointing into that list)"""
    er = np.asarray(sub.edge_row, dtype=np.uint32)
    ok = er != MISSING_EDGE
    rows, inv = np.unique(er[ok], return_inverse=True)
    er2 = er.copy()
    er2[ok] = inv.astype(np.uint32)
    return rows.astype(np.uint32), SnarlPaths(sub.snarl_path_off, sub.path_edge_off, er2)


def replay_deferred_on_device(shard_ctxs, col_ranges, paths: SnarlPaths, todo: np.ndarray, y, thresholds, replay_ctx,
                              allgather_dev=None) -> dict:
    """The snarls a sample-sharded solve deferred, finished WITHOUT leaving the device: every shard gathers its column
    slices of the rows those snarls use (stoat_gpu_gather_rows_dev), the slices are exchanged (`allgather_dev`: NCCL
    all-gather of one byte tensor per rank in a distributed run; None when `shard_ctxs` already holds every shard, as in
    a single-process run over several contexts), joined at their byte offsets into whole rows and handed to `replay_ctx`
    (a second context: the shard's own matrix and staging stay as they are) for the ordinary batch call."""
    import torch

    mi, mh, maf = thresholds
    sub = select_snarls(paths, todo)
    rows, sub = remap_rows(sub)
    dev = torch.device("cuda", replay_ctx.device)
    n_cols = sum(c1 - c0 for c0, c1 in col_ranges)
    full_pitch = (n_cols + 255) // 256 * 32
    parts = []
    for ctx in shard_ctxs:
        ptr, pitch = ctx.gather_rows_dev(rows)
        t = torch.as_tensor(_DevU8(ptr, rows.size * pitch), device=torch.device("cuda", ctx.device)).view(rows.size, pitch)
        parts.append(t)
    if allgather_dev is not None:
        parts = allgather_dev(parts[0])
    full = torch.zeros((rows.size, full_pitch), dtype=torch.uint8, device=dev)
    for (c0, c1), t in zip(col_ranges, parts):
        if c1 <= c0:
            continue
        nb = (c1 - c0 + 7) // 8
        full[:, c0 // 8:c0 // 8 + nb] = t[:, :nb].to(dev)
    torch.cuda.synchronize(dev)
    replay_ctx.set_matrix_pitched_dev(full.data_ptr(), int(rows.size), n_cols, full_pitch)
    replay_ctx.set_quantitative_phenotype(np.asarray(y, dtype=np.float64))
    return replay_ctx.quantitative_batch(sub, mi, mh, maf)


class _DevU8:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 3}


def run_quantitative_sample_sharded(ctx, bits, n_rows: int, n_cols: int, paths: SnarlPaths, y, rank: int, world: int,
                                    group=None, thresholds=(3, 5, 0.05), replay_ctx=None) -> dict:
    """One rank's part of a sample-sharded quantitative run (every rank returns the full result).

    moments of the local columns (K3M: per snarl 7 class counts + 3 phenotype sums) -> ONE all-reduce -> filters and OLS
    on the sums (K4).  Snarls that need more than the two-path sums (other shapes, rank-deficient or near-perfect
    fits: a few per cent) come back deferred; the column slices of their rows are all-gathered over NCCL (device buffers,
    no pickling, no host copy of the matrix), joined on the device and replayed through the ordinary batch call on a second
    context, replicated on every rank.  `ctx` keeps its shard (matrix, phenotype slice, staging)."""
    import torch.distributed as dist

    ranges = plan_column_shards(n_cols, world)
    c0, c1 = ranges[rank]
    if c1 <= c0:
        raise ValueError("sample-sharded run needs at least 128 columns per rank")
    mi, mh, maf = thresholds
    yy = np.asarray(y, dtype=np.float64)
    mean, s1, s2, vmax = phenotype_summary(yy)
    local = split_columns(bits, n_rows, n_cols, c0, c1)
    ctx.set_matrix(local, n_rows, c1 - c0)
    ctx.set_quantitative_phenotype_shard(yy[c0 // 2:c1 // 2], mean, s1, s2, vmax)
    ctx.stage_paths(paths)
    ctx.quantitative_moments_staged(mi, mh, maf)
    if world > 1:
        allreduce_partials_inplace(ctx, group)
    ctx.quantitative_solve_staged(mi, mh, maf)
    res = ctx.quantitative_fetch()
    todo = deferred_snarls(res)
    if todo.size == 0:
        return res
    from .api import Context

    def allgather_dev(mine):
        """one byte tensor [rows, local pitch] per rank; pitches differ by rank, so pad to the widest"""
        import torch

        if world == 1:
            return [mine]
        w = torch.tensor([mine.shape[1]], dtype=torch.int64, device=mine.device)
        ws = [torch.zeros_like(w) for _ in range(world)]
        dist.all_gather(ws, w, group=group)
        wmax = max(int(v.item()) for v in ws)
        padded = torch.zeros((mine.shape[0], wmax), dtype=torch.uint8, device=mine.device)
        padded[:, :mine.shape[1]] = mine
        out = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(out, padded, group=group)
        return out

    own = replay_ctx is None
    rctx = Context(ctx.device) if own else replay_ctx
    try:
        replay = replay_deferred_on_device([ctx], ranges, paths, todo, yy, thresholds, rctx, allgather_dev)
    finally:
        if own:
            rctx.close()
    return merge_deferred(res, todo, replay, paths)


def run_quantitative_sample_sharded_local(ctxs, bits, n_rows: int, n_cols: int, paths: SnarlPaths, y,
                                          thresholds=(3, 5, 0.05), replay_ctx=None) -> dict:
    """The same run inside ONE process over len(ctxs) contexts (on one GPU or several): shard i holds the column range i.
    The all-reduce becomes a sum over the contexts' partial buffers, the exchange of deferred rows a list.  This is what a
    single-GPU test box can run; the arithmetic and every library call are those of the distributed driver."""
    import torch

    from .api import Context

    world = len(ctxs)
    ranges = plan_column_shards(n_cols, world)
    mi, mh, maf = thresholds
    yy = np.asarray(y, dtype=np.float64)
    mean, s1, s2, vmax = phenotype_summary(yy)
    views = []
    for ctx, (c0, c1) in zip(ctxs, ranges):
        if c1 <= c0:
            raise ValueError("sample-sharded run needs at least 128 columns per shard")
        ctx.set_matrix(split_columns(bits, n_rows, n_cols, c0, c1), n_rows, c1 - c0)
        ctx.set_quantitative_phenotype_shard(yy[c0 // 2:c1 // 2], mean, s1, s2, vmax)
        ctx.stage_paths(paths)
        ctx.quantitative_moments_staged(mi, mh, maf)
        pi, ni, pd, nd = ctx.quantitative_partials_dev()
        dev = torch.device("cuda", ctx.device)
        views.append((torch.as_tensor(_DevU32(pi, ni), device=dev), torch.as_tensor(_DevF64(pd, nd), device=dev)))
    if len({v[0].numel() for v in views}) != 1:
        raise ValueError("the shards use different numbers of haplotype slices: give every shard <= 12,288 haplotypes")
    tot_i = sum(v[0].to(views[0][0].device) for v in views)
    tot_d = sum(v[1].to(views[0][1].device) for v in views)
    for vi, vd in views:
        vi.copy_(tot_i.to(vi.device))
        vd.copy_(tot_d.to(vd.device))
    for ctx in ctxs:
        torch.cuda.synchronize(ctx.device)
    ctxs[0].quantitative_solve_staged(mi, mh, maf)
    res = ctxs[0].quantitative_fetch()
    todo = deferred_snarls(res)
    if todo.size == 0:
        return res
    own = replay_ctx is None
    rctx = Context(ctxs[0].device) if own else replay_ctx
    try:
        replay = replay_deferred_on_device(ctxs, ranges, paths, todo, yy, thresholds, rctx)
    finally:
        if own:
            rctx.close()
    return merge_deferred(res, todo, replay, paths)
