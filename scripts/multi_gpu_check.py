#!/usr/bin/env python3
"""One rank of a multi-GPU parity run (launched by torchrun, one process per GPU, NCCL over NVLink).

Both shardings of SURVEY §8e on a seeded problem; rank 0 compares with the oracle's single-process result:
  * snarl-sharded binary and quantitative (no data-path collective, host gather of the result arrays)
  * sample-sharded quantitative: K3M moments of the local columns, ONE all-reduce of the partial sums, K4 on the sums
  * sample-sharded binary: count pass on the local columns, ONE ncclAllReduce(sum) applied in place to the context's
    device count buffer, statistics on the summed counts — every rank must hold the full, bit-exact result.
Exit code 0 = parity.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H  # noqa: E402
import oracle  # noqa: E402
import stoat_b200  # noqa: E402
from stoat_b200 import shard, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    d = synth.numpy_dataset(3000, 640, seed=17, frac_missing_edge=0.02)  # 1280 haplotypes: 10 chunks of 128 columns
    paths, bits, n_rows, n_cols = d["paths"], d["bits"], d["n_rows"], d["n_cols"]
    P = oracle.Paths(paths.snarl_path_off, paths.path_edge_off, paths.edge_row)
    ok = True
    with stoat_b200.Context(local) as ctx:
        got = shard.run_snarl_sharded(ctx, "binary", bits, n_rows, n_cols, paths, d["is_case"], rank, world)
        if rank == 0:
            H.assert_binary_equal(got, oracle.binary_batch(bits, n_cols, d["is_case"], P))
            print(f"snarl-sharded binary x{world}: parity ok ({P.n_snarls} snarls)", flush=True)
        got = shard.run_snarl_sharded(ctx, "quantitative", bits, n_rows, n_cols, paths, d["y"], rank, world)
        if rank == 0:
            H.assert_quant_equal(got, oracle.quantitative_batch(bits, n_cols, d["y"], P))
            print(f"snarl-sharded quantitative x{world}: parity ok", flush=True)
        got = shard.run_binary_sample_sharded(ctx, bits, n_rows, n_cols, paths, d["is_case"], rank, world)
        try:
            H.assert_binary_equal(got, oracle.binary_batch(bits, n_cols, d["is_case"], P))  # on EVERY rank
        except AssertionError as e:
            print(f"rank {rank}: sample-sharded mismatch: {e}", flush=True)
            ok = False
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"sample-sharded binary x{world} (NCCL all-reduce of the count buffer): "
                  f"{'parity ok on every rank' if flag.item() else 'MISMATCH'}", flush=True)
        ok = bool(flag.item())
        # sample-sharded quantitative: local moments (K3M) -> NCCL all-reduce of the partial sums -> solve (K4); deferred
        # snarls replayed from the re-assembled rows
        got = shard.run_quantitative_sample_sharded(ctx, bits, n_rows, n_cols, paths, d["y"], rank, world)
        okq = True
        try:
            H.assert_quant_equal(got, oracle.quantitative_batch(bits, n_cols, d["y"], P))  # on EVERY rank
        except AssertionError as e:
            print(f"rank {rank}: sample-sharded quantitative mismatch: {e}", flush=True)
            okq = False
        flag = torch.tensor([1 if okq else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"sample-sharded quantitative x{world} (NCCL all-reduce of the class counts + phenotype sums): "
                  f"{'parity ok on every rank' if flag.item() else 'MISMATCH'}", flush=True)
        ok = ok and bool(flag.item())
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
