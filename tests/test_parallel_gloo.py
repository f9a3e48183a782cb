"""World-size-2 gloo tests (CPU) of the multi-GPU plumbing: sharding, ragged all-gather,
all-reduce of partial sums, and the prompt-sharded extraction protocol.  The CUDA kernels are
replaced by the numpy oracle here (tests may use it); on the GPU box the same protocol runs with
K1a/K1b (bench.py at --gpus > 1)."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from introspective_awareness_b200 import parallel


def test_sharding_helpers():
    assert parallel.shard_round_robin(list(range(32)), 3, 8) == [3, 11, 19, 27]
    assert [parallel.shard_contiguous(50, r, 8) for r in range(8)] == [
        (0, 7), (7, 14), (14, 20), (20, 26), (26, 32), (32, 38), (38, 44), (44, 50)]
    assert parallel.shard_contiguous(3, 5, 8) == (3, 3)  # more ranks than items: empty shard
    cover = []
    for r in range(5):
        lo, hi = parallel.shard_contiguous(42, r, 5)
        cover += list(range(lo, hi))
    assert cover == list(range(42))  # Gemma-2's 42 layers over 5 ranks: every layer exactly once
    # the sweep's deal: the same number of whole layers for every rank, the left-over ones shared batch by batch
    whole, shared = parallel.shard_whole_and_shared(list(range(42)), 8)
    assert [len(w) for w in whole] == [5] * 8 and shared == [40, 41] and whole[3] == [3, 11, 19, 27, 35]
    assert sorted(x for w in whole for x in w) + shared == list(range(42))
    assert parallel.shard_whole_and_shared([7], 8) == ([[] for _ in range(8)], [7])  # one layer on 8 GPUs: shared by all
    assert parallel.shard_whole_and_shared(list(range(32)), 8)[1] == [] and parallel.shard_whole_and_shared([1, 2], 1) == ([[1, 2]], [])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from oracle import kernels_np as K

    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    r, w, _ = parallel.init_from_env(backend="gloo")
    assert (r, w) == (rank, world) and parallel.world() == (rank, world)

    # ragged all-gather along dim 1 and rank-0 gather
    t = torch.full((2, rank + 1, 3), float(rank))
    g = parallel.allgather_rows(t, [1, 2], dim=1)
    assert g.shape == (2, 3, 3) and g[:, 0].eq(0).all() and g[:, 1:].eq(1).all()
    parts = parallel.gather_to_rank0(torch.tensor([rank * 10 + 1]))
    assert (parts is None) == (rank != 0)
    if rank == 0:
        assert [int(p) for p in parts] == [1, 11]

    # prompt-sharded extraction == single-rank extraction
    rng = np.random.default_rng(0)
    L, C, Nb, H = 3, 5, 101, 64
    conc = K.rd((rng.standard_normal((L, C, H)) * 3 + 5).astype(np.float32), "bf16")
    base = K.rd((rng.standard_normal((L, Nb, H)) * 3 + 5).astype(np.float32), "bf16")

    def capture_rows_and_sum(segment, lo, hi):
        src = conc if segment == "concept" else base
        rows = torch.from_numpy(src[:, lo:hi].copy())
        sums = torch.from_numpy(np.stack([K.column_sum(src[l, lo:hi]) for l in range(L)]))
        return rows, sums

    def finalize(rows, base_sum, n_base, mode, normalize):
        outs = [K.meandiff(rows[l].numpy(), None, "raw" if base_sum is None else mode, "bf16", normalize,
                           base_sum=None if base_sum is None else base_sum[l].numpy(), n_base=n_base)
                for l in range(L)]
        return np.stack([o[1] for o in outs])

    got = parallel.extract_vectors_sharded(capture_rows_and_sum, finalize, C, Nb, "baseline", False)
    want = np.stack([K.meandiff(conc[l], base[l], "baseline", "bf16")[1] for l in range(L)])
    rel = np.linalg.norm(got - want, axis=-1) / np.linalg.norm(want, axis=-1)
    assert rel.max() <= 1e-3, rel.max()
    # only the fp32 summation order differs (sum of two partial sums): almost every element identical
    assert (got != want).mean() <= 1e-2
    raw = parallel.extract_vectors_sharded(capture_rows_and_sum, finalize, C, Nb, "no_baseline", False)
    assert np.array_equal(raw, conc)
    np.save(os.path.join(out_dir, f"vec{rank}.npy"), got)

    # the sweep through the product API on two ranks: layer fractions dealt round-robin, every rank writes the
    # files of its own configurations, result rows gathered to rank 0
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_reference_sweep_dropin import ScriptedModel
    from introspective_awareness_b200 import sweep, vector_utils as vu

    model = ScriptedModel(chat=True)
    res = sweep.run_sweep(model, ["Dust", "Trees"], vu.get_baseline_words(4), [0.2, 0.5, 0.8], [2.0, 8.0], n_trials=4,
                          max_new_tokens=5, temperature=0.0, batch_size=8, output_dir=os.path.join(out_dir, "sweep"))
    mine = {(f, a) for i, f in enumerate([0.2, 0.5, 0.8]) if i % world == rank for a in (2.0, 8.0)}
    assert set(res) == ({(f, a) for f in (0.2, 0.5, 0.8) for a in (2.0, 8.0)} if rank == 0 else mine)
    if rank == 0:
        torch.save(res, os.path.join(out_dir, "sweep_rank0.pt"))
    # fewer layers than ranks: the one layer's row batches are shared between the ranks, rank 0 puts them together
    res1 = sweep.run_sweep(ScriptedModel(chat=True), ["Dust", "Trees", "Oceans"], vu.get_baseline_words(4), [0.5], [2.0, 8.0], n_trials=5,
                           max_new_tokens=5, temperature=0.0, batch_size=4, output_dir=os.path.join(out_dir, "sweep1"))
    assert set(res1) == ({(0.5, 2.0), (0.5, 8.0)} if rank == 0 else set())
    if rank == 0:
        torch.save(res1, os.path.join(out_dir, "sweep1_rank0.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = np.load(tmp_path / "vec0.npy"), np.load(tmp_path / "vec1.npy")
    assert np.array_equal(a, b)  # every rank ends with the same vectors
    # the two-rank sweep gathered on rank 0 == the same sweep in one process; every configuration's files exist
    import json
    import sys

    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from test_reference_sweep_dropin import ScriptedModel
    from introspective_awareness_b200 import sweep, vector_utils as vu

    two = torch.load(tmp_path / "sweep_rank0.pt")
    one = sweep.run_sweep(ScriptedModel(chat=True), ["Dust", "Trees"], vu.get_baseline_words(4), [0.2, 0.5, 0.8], [2.0, 8.0],
                          n_trials=4, max_new_tokens=5, temperature=0.0, batch_size=8, output_dir=str(tmp_path / "single"))
    assert two == one
    assert list(two) == list(one)  # and in the caller's order
    for (f, a), rows in one.items():
        blob = json.load(open(tmp_path / "sweep" / "spy_model" / f"layer_{f:.2f}_strength_{a}" / "results.json"))
        assert blob["results"] == rows
    # 0.2 and 0.5 were whole layers of rank 0 / rank 1, 0.8 was shared batch by batch; a single shared layer likewise
    shared = torch.load(tmp_path / "sweep1_rank0.pt")
    alone = sweep.run_sweep(ScriptedModel(chat=True), ["Dust", "Trees", "Oceans"], vu.get_baseline_words(4), [0.5], [2.0, 8.0],
                            n_trials=5, max_new_tokens=5, temperature=0.0, batch_size=4, output_dir=str(tmp_path / "single1"))
    assert shared == alone and len(alone[(0.5, 2.0)]) == 3 * (2 + 3 + 2)
    for (f, a), rows in alone.items():
        for root in ("sweep1", "single1"):
            blob = json.load(open(tmp_path / root / "spy_model" / f"layer_{f:.2f}_strength_{a}" / "results.json"))
            assert blob["results"] == rows
        assert (open(tmp_path / "sweep1" / "spy_model" / f"layer_{f:.2f}_strength_{a}" / "results.csv").read()
                == open(tmp_path / "single1" / "spy_model" / f"layer_{f:.2f}_strength_{a}" / "results.csv").read())
