"""Host-side sharding logic (stoat_b200/shard.py) under torch.distributed `gloo`, world_size 2, on CPU.

The partition / remap / merge code is what the multi-GPU runs use; here the per-rank compute is the oracle
(tests may use it as a stand-in), so these tests pin the plumbing: a sharded run must reproduce the
single-process result exactly."""
from __future__ import annotations

import os
import socket
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import helpers  # noqa: E402

from stoat_b200 import shard  # noqa: E402
from stoat_b200.api import SnarlPaths  # noqa: E402


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(seed=5, n_snarls=60, n_samples=160, unaligned=False):
    rng = np.random.default_rng(seed)
    if unaligned:
        n_samples = 75  # 150 columns: rows are not byte aligned in the reference layout
    dense, snarls = helpers.partition_problem(rng, n_snarls, n_samples, max_paths=4)
    # a snarl sharing a row with its neighbour, an empty path list entry and a missing edge
    snarls[3][0] = [snarls[2][0][0]]
    snarls[7].append([])
    snarls[9][1] = [helpers.MISSING]
    spo, peo, er = helpers.csr_from_lists(snarls)
    is_case = (rng.random(n_samples) < 0.5).astype(np.uint8)
    bits = np.packbits(dense.reshape(-1), bitorder="little")
    return bits, dense.shape[0], dense.shape[1], SnarlPaths(spo, peo, er), is_case


def test_plans_cover_and_align():
    bits, n_rows, n_cols, paths, _ = _problem()
    for world in (1, 2, 3, 8, 100):
        plan = shard.plan_snarl_shards(paths, world)
        assert len(plan) == world and plan[0][0] == 0 and plan[-1][1] == paths.n_snarls
        assert all(a[1] == b[0] for a, b in zip(plan, plan[1:])) and all(a <= b for a, b in plan)
    for n, world in ((10000, 8), (320, 2), (130, 4), (128, 2), (2, 2)):
        plan = shard.plan_column_shards(n, world)
        assert plan[0][0] == 0 and plan[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(plan, plan[1:]))
        assert all(c0 % 128 == 0 for c0, _ in plan if c0 < n)


@pytest.mark.parametrize("unaligned", [False, True])
def test_extract_shard_is_self_contained(oracle, unaligned):
    bits, n_rows, n_cols, paths, is_case = _problem(unaligned=unaligned)
    full = oracle.binary_batch(bits, n_cols, is_case, oracle.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row))
    parts = []
    for s0, s1 in shard.plan_snarl_shards(paths, 3):
        sh = shard.extract_snarl_shard(bits, n_rows, n_cols, paths, s0, s1)
        assert sh.n_rows == sh.rows.size and (sh.n_rows == 0 or sh.rows.max() < n_rows)
        lp = oracle.Paths(sh.paths.snarl_path_off, sh.paths.path_edge_off, sh.paths.edge_row)
        r = oracle.binary_batch(sh.bits if sh.n_rows else np.zeros(1, np.uint8), n_cols, is_case, lp)
        r.pop("seconds")
        parts.append(r)
    merged = shard.merge_snarl_results(parts)
    for k in merged:
        np.testing.assert_array_equal(merged[k], full[k], err_msg=k)


def _worker(rank, world, port, mode, out_path):
    import torch
    import torch.distributed as dist

    import oracle as O

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bits, n_rows, n_cols, paths, is_case = _problem()
        if mode == "snarl":
            s0, s1 = shard.plan_snarl_shards(paths, world)[rank]
            sh = shard.extract_snarl_shard(bits, n_rows, n_cols, paths, s0, s1)
            lp = O.Paths(sh.paths.snarl_path_off, sh.paths.path_edge_off, sh.paths.edge_row)
            part = O.binary_batch(sh.bits if sh.n_rows else np.zeros(1, np.uint8), n_cols, is_case, lp)
            part.pop("seconds")
            gathered = [None] * world if rank == 0 else None
            dist.gather_object(part, gathered, dst=0)  # host merge only: no data-path collective
            if rank == 0:
                np.savez(out_path, **shard.merge_snarl_results(gathered))
        elif mode == "deferred":
            # the replay step of the sample-sharded quantitative run: whole rows of a few snarls rebuilt from every
            # rank's column slice (all_gather_object), edge_row remapped to the compact matrix
            ranges = shard.plan_column_shards(n_cols, world)
            c0, c1 = ranges[rank]
            local = shard.split_columns(bits, n_rows, n_cols, c0, c1)
            todo = np.array([2, 3, 7, 9, 31, 58])

            def allgather(obj):
                out = [None] * world
                dist.all_gather_object(out, obj)
                return out

            fb, fr, sub = shard.gather_deferred_rows(local, n_rows, c1 - c0, paths, todo, ranges, allgather)
            if rank == 1:
                np.savez(out_path, bits=fb, n_rows=fr, spo=sub.snarl_path_off, peo=sub.path_edge_off, er=sub.edge_row)
        else:
            c0, c1 = shard.plan_column_shards(n_cols, world)[rank]
            local = shard.split_columns(bits, n_rows, n_cols, c0, c1)
            lp = O.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row)
            r = O.binary_batch(local, c1 - c0, is_case[c0 // 2:c1 // 2], lp)
            counts = torch.from_numpy(np.concatenate([r["g0"], r["g1"]]).astype(np.int32))
            dist.all_reduce(counts, op=dist.ReduceOp.SUM)  # the ONE exchange of the sample-sharded mode
            if rank == 0:
                P = paths.n_paths
                np.savez(out_path, g0=counts[:P].numpy().astype(np.uint32), g1=counts[P:].numpy().astype(np.uint32))
    finally:
        dist.destroy_process_group()


def _spawn(mode, tmp_path):
    import torch.multiprocessing as mp

    out = str(tmp_path / f"{mode}.npz")
    mp.spawn(_worker, args=(2, _free_port(), mode, out), nprocs=2, join=True)
    return np.load(out)


def test_snarl_sharded_world2_matches_single_process(oracle, tmp_path):
    got = _spawn("snarl", tmp_path)
    bits, n_rows, n_cols, paths, is_case = _problem()
    full = oracle.binary_batch(bits, n_cols, is_case, oracle.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row))
    for k in got.files:
        np.testing.assert_array_equal(got[k], full[k], err_msg=k)


def test_sample_sharded_world2_counts_allreduce(oracle, tmp_path):
    got = _spawn("sample", tmp_path)
    bits, n_rows, n_cols, paths, is_case = _problem()
    full = oracle.binary_batch(bits, n_cols, is_case, oracle.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row))
    np.testing.assert_array_equal(got["g0"], full["g0"])
    np.testing.assert_array_equal(got["g1"], full["g1"])
    # the statistics pass sees exactly the single-process counts, hence the same tables / p-values
    spo = paths.snarl_path_off.astype(np.int64)
    for s in range(paths.n_snarls):
        g0 = got["g0"][spo[s]:spo[s + 1]].astype(np.uint64)
        g1 = got["g1"][spo[s]:spo[s + 1]].astype(np.uint64)
        assert int((g0 + g1).sum()) == int(full["total_sum"][s])


def test_deferred_rows_reassembled_world2(oracle, tmp_path):
    """Sample-sharded quantitative run, replay step: the compact full-width matrix + sub-batch built from two ranks'
    column slices gives, snarl by snarl, what the whole matrix gives (oracle as the compute)."""
    got = _spawn("deferred", tmp_path)
    bits, n_rows, n_cols, paths, _ = _problem()
    todo = np.array([2, 3, 7, 9, 31, 58])
    rng = np.random.default_rng(3)
    y = rng.normal(size=n_cols // 2)
    full = oracle.quantitative_batch(bits, n_cols, y, oracle.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row),
                                     min_ind=2, min_hap=2, maf=0.01)
    part = oracle.quantitative_batch(got["bits"], n_cols, y, oracle.Paths(got["spo"], got["peo"], got["er"]),
                                     min_ind=2, min_hap=2, maf=0.01)
    for k in ("filtered", "n_used", "code"):
        np.testing.assert_array_equal(part[k], full[k][todo], err_msg=k)
    for k in ("p_value", "beta", "se", "r2"):
        np.testing.assert_array_equal(np.nan_to_num(part[k], nan=-7.0), np.nan_to_num(full[k][todo], nan=-7.0), err_msg=k)
    merged = shard.merge_deferred({k: np.zeros_like(v) for k, v in full.items() if k != "seconds"}, todo,
                                  {k: v for k, v in part.items() if k != "seconds"}, paths)
    spo = paths.snarl_path_off.astype(np.int64)
    for s in todo:
        np.testing.assert_array_equal(merged["allele_paths"][spo[s]:spo[s + 1]], full["allele_paths"][spo[s]:spo[s + 1]])
        assert merged["n_used"][s] == full["n_used"][s]


def test_phenotype_summary_matches_library_definition():
    rng = np.random.default_rng(8)
    y = rng.normal(size=777) * 5 + 100
    mean, s1, s2, vmax = shard.phenotype_summary(y)
    assert abs(mean - y.mean()) < 1e-12 and abs(s1) < 1e-9
    assert abs(s2 - ((y - mean) ** 2).sum()) < 1e-9 * s2 and vmax == np.abs(y - mean).max()
    sel = shard.select_snarls(_problem()[3], np.array([5, 1]))
    full = _problem()[3]
    spo, peo = full.snarl_path_off.astype(np.int64), full.path_edge_off.astype(np.int64)
    assert sel.n_snarls == 2 and sel.n_paths == (spo[6] - spo[5]) + (spo[2] - spo[1])
    np.testing.assert_array_equal(sel.edge_row[: peo[spo[6]] - peo[spo[5]]], full.edge_row[peo[spo[5]]:peo[spo[6]]])


def test_select_snarls_vectorised_equals_loops():
    from stoat_b200 import shard, synth

    d = synth.numpy_dataset(500, 20, seed=3, frac_missing_edge=0.05)
    p = d["paths"]
    idx = np.array([7, 3, 499, 0, 250, 251])
    sub = shard.select_snarls(p, idx)
    spo, peo = p.snarl_path_off.astype(np.int64), p.path_edge_off.astype(np.int64)
    want_er, want_peo, want_spo = [], [0], [0]
    for s_ in idx:
        for q in range(spo[s_], spo[s_ + 1]):
            want_er.extend(p.edge_row[peo[q]:peo[q + 1]].tolist())
            want_peo.append(len(want_er))
        want_spo.append(len(want_peo) - 1)
    np.testing.assert_array_equal(sub.snarl_path_off, np.array(want_spo, np.uint64))
    np.testing.assert_array_equal(sub.path_edge_off, np.array(want_peo, np.uint64))
    np.testing.assert_array_equal(sub.edge_row, np.array(want_er, np.uint32))


def test_merge_deferred_vectorised():
    from stoat_b200 import shard, synth

    d = synth.numpy_dataset(60, 20, seed=4)
    p = d["paths"]
    spo = p.snarl_path_off.astype(np.int64)
    res = dict(allele_paths=np.zeros(p.n_paths, np.uint32), p_value=np.zeros(p.n_snarls), code=np.full(p.n_snarls, 7, np.uint8))
    todo = np.array([3, 10, 59])
    n_paths = int((spo[todo + 1] - spo[todo]).sum())
    replay = dict(allele_paths=np.arange(1, n_paths + 1, dtype=np.uint32), p_value=np.array([0.1, 0.2, 0.3]), code=np.zeros(3, np.uint8))
    out = shard.merge_deferred(res, todo, replay, p)
    assert out["p_value"][10] == 0.2 and out["code"][59] == 0 and out["code"][0] == 7
    pos = 0
    for s_ in todo:
        n = int(spo[s_ + 1] - spo[s_])
        np.testing.assert_array_equal(out["allele_paths"][spo[s_]:spo[s_] + n], replay["allele_paths"][pos:pos + n])
        pos += n
    assert out["allele_paths"].sum() == replay["allele_paths"].sum()
