"""Synthetic snarls x haplotypes data sets (SURVEY §8d, configs C1'-C5).

Structure of a data set (decided on the host with numpy, seed 42 by default):
  * every snarl has nodes a,b,c,d: paths >a>b>d and >a>c>d  => 4 distinct matrix rows, 2 checked edges each;
  * `frac_three` of the snarls get a third path >a>e>d (2 more rows) => 2xN chi2, no Fisher;
  * `frac_node0` get, instead of a normal alt path, a path whose edges all touch node 0 ('*', complex) so that
    it has NO checked edge and matches every haplotype (src/snarl_analyzer.cpp:328-330, 343-354);
  * `frac_missing_edge` get an alt path with an edge that is absent from the matrix (=> matches nobody, :334-336);
  * allele frequency f ~ U(0,1) per snarl, haplotype ~ Bernoulli(f) (tests/simu/scripts/simulate_test_dataset.py:146-150);
    `frac_marker` snarls shift the case group's frequency by U(0.01, 0.5) (:238-262), capped so chi2 stays < 200;
  * `frac_missing_hap` of the haplotypes of each snarl are on neither path.
  * phenotype: first half of the samples controls, second half cases (:53-62); quantitative y ~ N(0,1) plus an
    effect on marker snarl 0.

`numpy_dataset` builds everything on the host (tests, CPU baseline samples); `torch_dataset` draws the
haplotype bits on the GPU in chunks directly into the reference's matrix layout (bench, n_cols % 8 == 0).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .api import SnarlPaths

MISSING_EDGE = 0xFFFFFFFF


@dataclass
class Structure:
    n_snarls: int
    kind: np.ndarray       # uint8 [S]: 0 biallelic, 1 three paths, 2 node-0 alt path, 3 missing-edge alt path
    n_rows_of: np.ndarray  # int64 [S] distinct rows of each snarl
    row0: np.ndarray       # int64 [S] first row
    freq: np.ndarray       # float64 [S]
    shift: np.ndarray      # float64 [S] case-group shift (0 for non-markers)
    paths: SnarlPaths
    n_rows: int


def make_structure(n_snarls: int, seed: int = 42, frac_three=0.05, frac_node0=0.02, frac_missing_edge=0.0,
                   frac_marker=0.10) -> Structure:
    rng = np.random.default_rng(seed)
    S = n_snarls
    u = rng.random(S)
    kind = np.zeros(S, np.uint8)
    kind[u < frac_three] = 1
    kind[(u >= frac_three) & (u < frac_three + frac_node0)] = 2
    kind[(u >= frac_three + frac_node0) & (u < frac_three + frac_node0 + frac_missing_edge)] = 3
    n_rows_of = np.where(kind == 1, 6, np.where(kind == 2, 2, 4)).astype(np.int64)
    row0 = np.concatenate([[0], np.cumsum(n_rows_of)[:-1]]).astype(np.int64) if S else np.zeros(0, np.int64)
    n_rows = int(n_rows_of.sum())
    freq = rng.random(S)
    marker = rng.random(S) < frac_marker
    shift = np.where(marker, rng.uniform(0.01, 0.5, S), 0.0)
    # keep effect sizes where the reference's 50-digit subtraction is still defined (chi2 < 200)
    shift = np.minimum(shift, 0.12)
    shift = np.where(freq + shift > 0.98, 0.0, shift)

    n_paths_of = np.where(kind == 1, 3, 2).astype(np.int64)
    snarl_path_off = np.concatenate([[0], np.cumsum(n_paths_of)]).astype(np.uint64)
    P = int(snarl_path_off[-1])
    # edges per path: 2, except the node-0 alt path (0 checked edges)
    path_snarl = np.repeat(np.arange(S), n_paths_of)
    path_idx = np.arange(P) - np.repeat(snarl_path_off[:-1].astype(np.int64), n_paths_of)
    n_edges_of = np.full(P, 2, np.int64)
    node0_alt = (kind[path_snarl] == 2) & (path_idx == 1)
    n_edges_of[node0_alt] = 0
    path_edge_off = np.concatenate([[0], np.cumsum(n_edges_of)]).astype(np.uint64)
    E = int(path_edge_off[-1])
    edge_path = np.repeat(np.arange(P), n_edges_of)
    edge_k = np.arange(E) - np.repeat(path_edge_off[:-1].astype(np.int64), n_edges_of)
    edge_row = (row0[path_snarl[edge_path]] + 2 * path_idx[edge_path] + edge_k).astype(np.int64)
    miss = (kind[path_snarl[edge_path]] == 3) & (path_idx[edge_path] == 1) & (edge_k == 1)
    edge_row[miss] = MISSING_EDGE
    paths = SnarlPaths(snarl_path_off, path_edge_off, edge_row.astype(np.uint32))
    return Structure(S, kind, n_rows_of, row0, freq, shift, paths, n_rows)


def phenotypes(n_samples: int, seed: int = 42):
    rng = np.random.default_rng(seed + 1)
    is_case = np.zeros(n_samples, np.uint8)
    is_case[n_samples // 2:] = 1
    y = rng.normal(size=n_samples)
    return is_case, y


def _allele_probabilities(st: Structure, s0: int, s1: int, is_case_hap: np.ndarray):
    """P(alt) per (snarl, haplotype) for snarls [s0, s1)."""
    f = st.freq[s0:s1, None]
    sh = st.shift[s0:s1, None]
    return f + sh * is_case_hap[None, :]


def numpy_dataset(n_snarls: int, n_samples: int, seed: int = 42, frac_missing_hap=0.01, **kw):
    """-> dict(bits, n_rows, n_cols, paths, is_case, y, structure).  Host only; sized for tests."""
    st = make_structure(n_snarls, seed, **kw)
    H = 2 * n_samples
    is_case, y = phenotypes(n_samples, seed)
    is_case_hap = np.repeat(is_case, 2).astype(np.float64)
    rng = np.random.default_rng(seed + 2)
    dense = np.zeros((st.n_rows, H), dtype=bool)
    step = max(1, 2_000_000 // max(H, 1))
    for s0 in range(0, n_snarls, step):
        s1 = min(n_snarls, s0 + step)
        p_alt = _allele_probabilities(st, s0, s1, is_case_hap)
        u = rng.random((s1 - s0, H))
        present = rng.random((s1 - s0, H)) >= frac_missing_hap
        alt = u < p_alt
        third = np.zeros_like(alt)
        k = st.kind[s0:s1]
        three = k == 1
        if three.any():
            # split the alt class in two for three-path snarls
            third[three] = alt[three] & (rng.random((int(three.sum()), H)) < 0.5)
        for i in range(s1 - s0):
            r0 = st.row0[s0 + i]
            ref_m = present[i] & ~alt[i]
            alt_m = present[i] & alt[i] & ~third[i]
            dense[r0] = ref_m
            dense[r0 + 1] = ref_m
            if k[i] == 2:
                continue  # node-0 alt path: no rows of its own
            dense[r0 + 2] = alt_m
            dense[r0 + 3] = alt_m
            if k[i] == 1:
                th_m = present[i] & third[i]
                dense[r0 + 4] = th_m
                dense[r0 + 5] = th_m
    bits = np.packbits(dense.reshape(-1), bitorder="little")
    # quantitative effect on the first marker-free snarl's alt dosage keeps some p-values small
    if n_snarls:
        dos = dense[st.row0[0] + (2 if st.kind[0] != 2 else 0)].reshape(n_samples, 2).sum(1)
        y = y + 0.5 * dos
    return dict(bits=bits, n_rows=st.n_rows, n_cols=H, paths=st.paths, is_case=is_case, y=y, structure=st,
                dense=dense)


def torch_dataset(n_snarls: int, n_samples: int, device, seed: int = 42, frac_missing_hap=0.01,
                  chunk_snarls: int = 0, cols: tuple | None = None, **kw):
    """Draw the matrix on the GPU (Philox, torch CUDA generator) straight into the reference's byte layout.

    -> dict(bits_dev (torch.uint8 [n_rows * n_cols / 8]), n_rows, n_cols, paths, is_case, y, structure)
    Requires n_cols % 8 == 0 (then rows are byte aligned in the reference layout).

    `cols` = (c0, c1): only the haplotype columns [c0, c1) of the cohort — the slice one GPU of a sample-sharded run
    holds (c0, c1 even).  The snarl structure and the phenotype split stay those of the whole cohort of
    `n_samples`; the draws of a slice come from their own Philox stream (seed, c0), so the ranks' slices side by side are
    one valid cohort.  n_cols / is_case / y of the result describe the slice.
    """
    import torch

    st = make_structure(n_snarls, seed, **kw)
    is_case, y = phenotypes(n_samples, seed)
    gen = torch.Generator(device=device)
    gen.manual_seed(seed)
    if cols is not None:
        c0, c1 = int(cols[0]), int(cols[1])
        if c0 % 2 or c1 % 2 or not (0 <= c0 < c1 <= 2 * n_samples):
            raise ValueError("cols must be an even, non-empty sub-range of the haplotype columns")
        is_case, y = is_case[c0 // 2:c1 // 2], y[c0 // 2:c1 // 2]
        n_samples = (c1 - c0) // 2
        gen.manual_seed(seed * 1_000_003 + c0 + 1)
    H = 2 * n_samples
    if H % 8:
        raise ValueError("torch_dataset needs n_cols % 8 == 0")
    if chunk_snarls <= 0:  # ~0.16e9 cells per chunk: three float32 draws of that size stay below 2 GB
        chunk_snarls = max(256, min(16384, (1 << 27) // max(H, 1) // 256 * 256))
    row_bytes = H // 8
    bits = torch.empty(st.n_rows * row_bytes, dtype=torch.uint8, device=device)
    out = bits.view(st.n_rows, row_bytes)
    case_hap = torch.from_numpy(np.repeat(is_case, 2).astype(np.float32)).to(device)
    weights = (2 ** torch.arange(8, device=device, dtype=torch.int32)).view(1, 1, 8)

    def pack(mask):  # bool [n, H] -> uint8 [n, H/8], LSB-first
        return (mask.view(mask.shape[0], row_bytes, 8).to(torch.int32) * weights).sum(-1).to(torch.uint8)

    for s0 in range(0, n_snarls, chunk_snarls):
        s1 = min(n_snarls, s0 + chunk_snarls)
        n = s1 - s0
        f = torch.from_numpy(st.freq[s0:s1].astype(np.float32)).to(device).view(n, 1)
        sh = torch.from_numpy(st.shift[s0:s1].astype(np.float32)).to(device).view(n, 1)
        kind = torch.from_numpy(st.kind[s0:s1].astype(np.int64)).to(device)
        p_alt = f + sh * case_hap.view(1, H)
        alt = torch.rand((n, H), device=device, generator=gen) < p_alt
        present = torch.rand((n, H), device=device, generator=gen) >= frac_missing_hap
        third = alt & (torch.rand((n, H), device=device, generator=gen) < 0.5) & (kind == 1).view(n, 1)
        ref_b = pack(present & ~alt)
        alt_b = pack(present & alt & ~third)
        thr_b = pack(present & third)
        rows6 = torch.stack([ref_b, ref_b, alt_b, alt_b, thr_b, thr_b], dim=1)  # [n, 6, row_bytes]
        n_rows_of = torch.from_numpy(st.n_rows_of[s0:s1]).to(device)
        keep = torch.arange(6, device=device).view(1, 6) < n_rows_of.view(n, 1)
        r0 = int(st.row0[s0])
        r1 = int(st.row0[s1 - 1] + st.n_rows_of[s1 - 1])
        out[r0:r1] = rows6[keep]
        del alt, present, third, ref_b, alt_b, thr_b, rows6
    return dict(bits_dev=bits, n_rows=st.n_rows, n_cols=H, paths=st.paths, is_case=is_case, y=y, structure=st)
