"""Two-rank NCCL test of the multi-GPU path through the product API (skipped on a box with fewer than two GPUs):
prompt-sharded extraction (K2-tapped forwards on each rank's slice of the prompts, K1a, NCCL all-reduce of the fp32
baseline sums and all-gather of the concept rows, K1b) against the un-sharded extraction on the same GPU, and
`sweep.run_sweep` as a collective call against the same sweep in one process."""

import os
import socket

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from introspective_awareness_b200 import parallel, sweep, vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r, w_, local = parallel.init_from_env(backend="nccl")
    assert (r, w_) == (rank, world)
    dev = f"cuda:{local}"
    m = ModelWrapper.from_arch("tiny-llama", device=dev, dtype=torch.bfloat16, seed=13)  # the same weights on every rank
    m.model_name = "tiny/llama"
    V = m.model.config.vocab_size
    c_ids, c_mask, _ = synthetic_prompt_ids(7, 20, 8, V, seed=2)    # 7 and 11: ragged shards (4 + 3, 6 + 5)
    b_ids, b_mask, _ = synthetic_prompt_ids(11, 20, 8, V, seed=3)
    layers = [0, 2, 3]
    for method in ("baseline", "simple", "no_baseline"):
        bi, bm = (b_ids[:1], b_mask[:1]) if method == "simple" else (b_ids, b_mask)
        _, sharded, _ = m.extract_vectors_sharded_ids(c_ids, c_mask, None if method == "no_baseline" else bi,
                                                      None if method == "no_baseline" else bm, layers, method)
        _, whole, _ = m.engine.extract_vectors(c_ids, c_mask, bi, bm, layers, mode=method, want_f32=False, want_norm=False)
        rel = ((sharded.float() - whole.float()).norm(dim=-1) / whole.float().norm(dim=-1)).max().item()
        assert rel <= 1e-3, (method, rel)                       # north_star's bound
        assert (sharded != whole).float().mean().item() <= 1e-2, method   # only the fp32 summation order can differ
        gathered = [torch.empty_like(sharded) for _ in range(world)]
        torch.distributed.all_gather(gathered, sharded)
        assert all(torch.equal(g, sharded) for g in gathered), method     # every rank ends with the same vectors
    # the sweep as a collective call: default (every rank extracts its own layers from the whole prompt sets: the
    # one-process bits) and with prompt-sharded extraction
    for tag, sharded in (("two", False), ("two_sharded", True)):
        res = sweep.run_sweep(m, ["Dust", "Trees", "Oceans"], vu.get_baseline_words(6), [0.3, 0.6, 0.8], [2.0, 8.0], n_trials=4,
                              max_new_tokens=5, temperature=0.0, batch_size=8, output_dir=os.path.join(out_dir, tag),
                              sharded_extraction=sharded)
        if rank == 0:
            assert set(res) == {(f, a) for f in (0.3, 0.6, 0.8) for a in (2.0, 8.0)}
            torch.save(res, os.path.join(out_dir, tag + ".pt"))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_world_size_2_nccl(tmp_path):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    from introspective_awareness_b200 import sweep, vector_utils as vu
    from introspective_awareness_b200.model_utils import ModelWrapper

    two, two_sharded = torch.load(tmp_path / "two.pt"), torch.load(tmp_path / "two_sharded.pt")
    m = ModelWrapper.from_arch("tiny-llama", device="cuda:0", dtype=torch.bfloat16, seed=13)
    m.model_name = "tiny/llama"
    one = sweep.run_sweep(m, ["Dust", "Trees", "Oceans"], vu.get_baseline_words(6), [0.3, 0.6, 0.8], [2.0, 8.0], n_trials=4,
                          max_new_tokens=5, temperature=0.0, batch_size=8, output_dir=str(tmp_path / "one"))
    # 0.3 / 0.6 are whole layers of rank 0 / rank 1, 0.8 is left over and shared batch by batch (rank 0 puts it together)
    assert two == one  # default extraction: every rank works from the whole prompt sets, so the one-process results exactly
    for k in sorted(one):
        assert torch.equal(torch.load(tmp_path / "two" / "tiny_llama" / "vectors" / f"layer_{k[0]:.2f}" / "Trees.pt"),
                           torch.load(tmp_path / "one" / "tiny_llama" / "vectors" / f"layer_{k[0]:.2f}" / "Trees.pt"))
    # prompt-sharded extraction: the vectors may differ in the last bits (each rank's forward is a smaller GEMM problem;
    # fp32 summation order of the sharded mean), which can move a greedy token on a random-init model: everything else —
    # rows, order, numbering — is the same
    assert set(two_sharded) == set(one)
    same = sum(a["response"] == b["response"] for k in one for a, b in zip(one[k], two_sharded[k]))
    total = sum(len(v) for v in one.values())
    assert [[(r["concept"], r["trial"], r["trial_type"], r["strength"]) for r in one[k]] for k in sorted(one)] == \
           [[(r["concept"], r["trial"], r["trial_type"], r["strength"]) for r in two_sharded[k]] for k in sorted(two_sharded)]
    assert same >= 0.9 * total, (same, total)
