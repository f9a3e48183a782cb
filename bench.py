"""Headline benchmark: steered trials/sec on the Llama-3.1-8B architecture (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   # reference CPU path

A "step" is one pass of the hot path over one batch of synthetic input (BASELINE.json configs[1]):
extract 50 concept vectors against 100 baseline prompts at the steering layer (two forwards + K1),
then run 200 steered trials (50 concepts x 4 strengths, one batch, a different (concept, alpha, t0)
per row) for 32 greedy tokens each (prompt pass + 31 cached decode steps, K2 at the layer in each).
With N > 1 every rank is an independent replica steering its own layer of the sweep
(configs[2], weak scaling); extraction is then prompt-sharded with one all-reduce of the fp32
baseline sums and one all-gather of the concept rows, and the token ids are gathered to rank 0.

Prints ONE JSON line on rank 0.  `value` times the path with inputs resident in HBM; `e2e` times the
same work through the public ModelWrapper-level API with pinned HOST buffers (H2D of ids / masks /
vectors / row metadata and D2H of vectors and token ids inside the timed region).  `roofline`
is the fused inject+RMSNorm kernel's prompt-pass launch (the largest launch of ours in a step),
timed live with CUDA events on its stream.  `cpu_baseline` is the reference's CPU path (the
torch restatement in oracle/ref_hooks.py: HF generate + forward hook) on the host cores, on a
bounded sample.  Nothing here reads /root/reference.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

ARCH = "llama-3.1-8b"
N_CONCEPTS, N_BASE = 50, 100
STRENGTHS = [1.0, 2.0, 4.0, 8.0]
T_PROMPT, T_MIN, N_NEW = 128, 112, 32
EXTRACT_T, EXTRACT_TMIN = 24, 8
LAYER_FRACTION = 0.7
METRIC = "steered trials/sec (Llama-3.1-8B)"
UNIT = "trials/s"


def peak_hbm():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profiles(add_form: bool = False):
    """dram read+write bytes per launch of the prompt-pass K2, from the committed ncu summary of the form
    that was launched (plain, or with the residual add folded in)."""
    try:
        name = "k2_add_prefill_ncu_summary.json" if add_form else "k2_prefill_ncu_summary.json"
        d = json.load(open(os.path.join(ROOT, "profiles", name)))
        return float(d["dram_bytes_per_launch"])
    except Exception:
        return None


class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.path)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------

def make_workload(vocab: int, rank: int):
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    c_ids, c_mask, _ = synthetic_prompt_ids(N_CONCEPTS, EXTRACT_T, EXTRACT_TMIN, vocab, seed=2)
    b_ids, b_mask, _ = synthetic_prompt_ids(N_BASE, EXTRACT_T, EXTRACT_TMIN, vocab, seed=3)
    rows = N_CONCEPTS * len(STRENGTHS)
    ids, mask, lens = synthetic_prompt_ids(rows, T_PROMPT, T_MIN, vocab, seed=1 + 100 * rank)
    t0 = (lens - 28).tolist()  # "the token before Trial N" sits ~28 tokens from the end
    row_vec = [i % N_CONCEPTS for i in range(rows)]
    row_alpha = [STRENGTHS[i // N_CONCEPTS] for i in range(rows)]
    return dict(c_ids=c_ids, c_mask=c_mask, b_ids=b_ids, b_mask=b_mask, ids=ids, mask=mask, lens=lens, t0=t0,
                row_vec=row_vec, row_alpha=row_alpha, rows=rows)


def run_ours(args):
    from introspective_awareness_b200 import kernels, parallel
    from introspective_awareness_b200.engine import SteerSpec
    from introspective_awareness_b200.model_utils import ModelWrapper, get_layer_at_fraction
    from introspective_awareness_b200.synthetic import build_model
    from introspective_awareness_b200.stub_tokenizer import StubTokenizer

    rank, world, local = parallel.init_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (our arm) needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    arch = args.arch
    model = build_model(arch, dtype=torch.bfloat16, device=str(dev), seed=None)
    w = ModelWrapper(arch, device=str(dev), dtype=torch.bfloat16, model=model,
                     tokenizer=StubTokenizer(vocab_size=model.config.vocab_size), attention=args.attention)
    eng = w.engine
    n_layers = eng.n_layers
    base_layer = get_layer_at_fraction(w, LAYER_FRACTION)
    # rank r steers its own layer of the sweep (weak scaling; N=1 -> the single layer of configs[1])
    sweep = [(base_layer + r) % n_layers for r in range(world)] if world > 1 else [base_layer]
    my_layer = sweep[rank]
    wl = make_workload(model.config.vocab_size, rank)
    rows = wl["rows"]
    on = lambda t: t.to(dev)
    c_ids, c_mask, b_ids, b_mask = on(wl["c_ids"]), on(wl["c_mask"]), on(wl["b_ids"]), on(wl["b_mask"])
    ids_d, mask_d = on(wl["ids"]), on(wl["mask"])
    pads = T_PROMPT - wl["lens"]
    t0_abs = (torch.tensor(wl["t0"]) + pads).to(torch.int32)
    row_vec_d = torch.tensor(wl["row_vec"], dtype=torch.int32, device=dev)
    row_alpha_d = torch.tensor(wl["row_alpha"], dtype=torch.float32, device=dev)
    t0_d = t0_abs.to(dev)
    # pinned host copies for the end-to-end arm
    pin = lambda t: t.clone().pin_memory()
    h_c_ids, h_c_mask, h_b_ids, h_b_mask = pin(wl["c_ids"]), pin(wl["c_mask"]), pin(wl["b_ids"]), pin(wl["b_mask"])
    h_ids, h_mask = pin(wl["ids"]), pin(wl["mask"])

    clo, chi = parallel.shard_contiguous(N_CONCEPTS, rank, world)
    blo, bhi = parallel.shard_contiguous(N_BASE, rank, world)
    sweep_idx = sweep.index(my_layer)

    def extract_resident(ci=None, cm=None, bi=None, bm=None):
        """-> [50, H] device vectors for this rank's layer (inputs default to the HBM-resident copies)."""
        ci, cm, bi, bm = (c_ids, c_mask, b_ids, b_mask) if ci is None else (ci, cm, bi, bm)
        if world == 1:
            _, vr, _ = eng.extract_vectors(ci, cm, bi, bm, [my_layer], want_f32=False, want_norm=False)
            return vr[0]
        # prompt-sharded: this rank forwards its slice of both prompt sets, for every layer of this step
        uniq = sorted(set(sweep))
        cl = eng.capture_layers(ci[clo:chi], cm[clo:chi], uniq, slot=0)
        bl = eng.capture_layers(bi[blo:bhi], bm[blo:bhi], uniq, slot=1)
        crow, _ = kernels.gather_segsum(cl, want_rows=True)
        _, bsum = kernels.gather_segsum(bl, want_rows=False)
        parallel.allreduce_sum_(bsum)
        counts = [b - a for a, b in (parallel.shard_contiguous(N_CONCEPTS, q, world) for q in range(world))]
        crow = parallel.allgather_rows(crow, counts, dim=1)
        _, vr, _ = kernels.meandiff_finalize(crow, bsum, N_BASE, "baseline", False, want_f32=False)
        return vr[uniq.index(my_layer)]

    sess = eng.session(rows, T_PROMPT, N_NEW, padded=True, n_vec_cap=64, vec_dtype=torch.bfloat16)

    def step_resident():
        table = extract_resident()
        steer = SteerSpec(my_layer, table, row_vec_d, row_alpha_d, t0_d)
        out = sess.generate(ids_d, mask_d, steer, temperature=0.0, eos_token_id=None, pad_token_id=0)
        if world > 1:
            parallel.gather_to_rank0(out)
        return out

    def step_e2e():
        """Public API with host buffers: ids/masks from pinned host memory, vectors returned to the host
        (as the reference's extraction does), generation fed with host vectors, ids read back."""
        vec_host = extract_resident(h_c_ids, h_c_mask, h_b_ids, h_b_mask).cpu()
        out = w.steered_generate_ids(h_ids, h_mask, my_layer, vec_host, wl["row_vec"], wl["row_alpha"], wl["t0"], N_NEW,
                                     temperature=0.0)
        host_ids = out.cpu()
        if world > 1:
            parallel.gather_to_rank0(out)
        return host_ids

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    eng.k2_events = []
    launches_before = dict(eng.counters)
    if args.ncu_window:
        # profiler window = the last timed step only (ncu --profile-from-start off -c N)
        real_step, seen = step_resident, [0]

        def step_resident():  # noqa: F811
            seen[0] += 1
            if seen[0] == args.steps:
                torch.cuda.profiler.start()
            real_step()

    ms_res = timed(step_resident, args.steps)
    launches_after = dict(eng.counters)
    if args.ncu_window:
        torch.cuda.profiler.stop()
        step_resident = real_step
    k2_events = eng.k2_events
    eng.k2_events = None
    clocks = sampler.stop() if sampler else None
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    trials = rows * world * args.steps
    value = trials / (ms_res / 1e3)
    e2e_value = trials / (ms_e2e / 1e3)

    # roofline of the dominant launch of ours: K2 over the prompt pass [rows, T_PROMPT, H]
    k2_ms = [a.elapsed_time(b) for a, b, _ in k2_events]
    H = eng.hidden
    tail = eng.fuse_gemma_tail  # Gemma-2: the block-tail variant reads mlp_out as a fifth stream and a second weight
    # Llama / Qwen2 (engine-driven block): the launch is the residual-add form — the MLP output is a fifth stream
    fifth = tail or eng.native_block
    algo = rows * T_PROMPT * H * 2 * (5 if fifth else 4) + H * 2 * (2 if tail else 1)
    peak, peak_src = peak_hbm()
    k2_avg_ms = sum(k2_ms) / max(len(k2_ms), 1)
    achieved = algo / (k2_avg_ms * 1e6) if k2_ms else None
    # what this particular launch actually moves: only positions >= t0 read a table row and write h'
    steered_vecs = int((T_PROMPT - t0_abs.clamp(max=T_PROMPT)).sum())
    moved = (rows * T_PROMPT * (4 if fifth else 2) + steered_vecs * (1 if fifth else 2)) * H * 2 + H * 2

    h2d = sum(t.numel() * t.element_size() for t in (h_c_ids, h_c_mask, h_b_ids, h_b_mask, h_ids, h_mask)) \
        + N_CONCEPTS * H * 2 + rows * (4 + 4 + 4)
    d2h = N_CONCEPTS * H * 2 + rows * N_NEW * 8
    # our kernels launched inside the value-timed region on this rank (counted by the engine: eager
    # launches directly, graph replays by the launches recorded at capture)
    launches_by_kernel = {k: launches_after[k] - launches_before[k] for k in launches_before}
    if world > 1:
        launches_by_kernel["k1_launches"] += args.steps * 4  # 2x gather_segsum + finalize + norm, called outside the engine
    launches = sum(launches_by_kernel.values())

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:  # reported at N=1 only
        cpu = cpu_baseline_port(model, my_layer, wl, sample_rows=args.cpu_rows)

    if rank == 0:
        line = {
            "metric": METRIC if arch == ARCH else f"steered trials/sec ({arch})", "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (random-init weights, stub-tokenised uniform ids)",
            "config": {
                "workload": f"BASELINE configs[1]: {arch} arch, {N_CONCEPTS} concepts x {len(STRENGTHS)} strengths at one "
                            f"layer, one {rows}-row steered batch per GPU, prompt {T_PROMPT} (real {T_MIN}..{T_PROMPT}, "
                            f"left-padded), {N_NEW} greedy tokens, extraction {N_CONCEPTS}+{N_BASE} prompts per step",
                "layers": sweep, "rows_per_gpu_per_step": rows, "new_tokens": N_NEW, "attention": eng.attention,
                "engine_block": "native" if eng.native_block else "hf",
                "single_gemm_projections": {f"{k}{list(shape)}": ok for (k, shape), ok in sorted(eng._fused_checked.items())},
                "parallelism": "independent replicas, layer axis round-robin" if world > 1 else "single replica",
                "l2": "inputs larger than L2: each step streams 16 GB of weights per decode step and ~1 GB of KV/activations",
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "gpu_launches_by_kernel": launches_by_kernel,
            "roofline": {"kernel": ("ia_add_inject_rmsnorm (ia::k2_tma_kernel, ADD form: residual add + inject + RMSNorm, cp.async.bulk "
                                    "staged), prompt-pass launch" if eng.native_block else
                                    "ia_gemma2_tail_inject_rmsnorm (ia::k2_tma_kernel, tail form: post-FFN norm + residual add + inject + "
                                    "next RMSNorm, cp.async.bulk staged), prompt-pass launch" if tail else
                                    "ia_inject_rmsnorm (ia::k2_tma_kernel: fused inject+RMSNorm, cp.async.bulk staged), prompt-pass launch"),
                         "bound": "hbm",
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if achieved else None,
                         "traffic": traffic_from_profiles(eng.native_block), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo, "us_per_launch": k2_avg_ms * 1e3,
                         "steered_fraction_of_vectors": steered_vecs / (rows * T_PROMPT),
                         "bytes_moved_per_launch": moved,
                         "frac_of_peak_on_bytes_moved": (moved / (k2_avg_ms * 1e6) / peak) if k2_ms else None,
                         "launches_timed": len(k2_ms),
                         "share_of_step": (k2_avg_ms * len(k2_ms) / ms_res) if k2_ms else None,
                         "streams_per_vector": 5 if fifth else 4,
                         "note": "achieved uses SURVEY 8(d)'s algorithmic bytes: residual in, table row in, h' out, y out for "
                                 "every vector, plus the branch output in when the layer's residual add is folded into the "
                                 "launch (streams_per_vector). In this launch only positions >= t0 "
                                 "(steered_fraction_of_vectors) read a table row, so fewer bytes actually move "
                                 "(bytes_moved_per_launch, frac_of_peak_on_bytes_moved); the branch output was also just "
                                 "written by the preceding GEMM (partly L2-resident). Cold-cache, every-vector-steered "
                                 "figures for the same shape: tools/kbench.py -> profiles/r1_kbench_ladder.jsonl; `traffic` "
                                 "is the DRAM byte count per launch from the ncu --set full capture of the launched form "
                                 "(profiles/k2_add_prefill_ncu_summary.json with the residual add, else "
                                 "profiles/k2_prefill_ncu_summary.json)"},
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# reference CPU path (oracle port)
# ---------------------------------------------------------------------------------------------

def _cpu_model_from(gpu_model):
    import copy

    m = copy.deepcopy(gpu_model).to("cpu")
    return m.eval()


def _fast_cpu_model(arch):
    """Random-init CPU model without HF's slow CPU initialisers (values only need a sane scale)."""
    from transformers import AutoModelForCausalLM

    from introspective_awareness_b200.synthetic import make_config

    cfg = make_config(arch)
    cfg._attn_implementation = "sdpa"
    with torch.device("meta"):
        m = AutoModelForCausalLM.from_config(cfg, attn_implementation="sdpa")
    m = m.to(torch.bfloat16).to_empty(device="cpu")
    with torch.no_grad():
        for name, p in m.named_parameters():
            if "norm" in name and p.dim() == 1:
                p.fill_(1.0)
            else:
                p.uniform_(-0.0346, 0.0346)
        for mod in m.modules():  # non-persistent buffers (rotary inv_freq) were dropped by to_empty
            if hasattr(mod, "rope_init_fn") and hasattr(mod, "inv_freq"):
                inv, _ = mod.rope_init_fn(mod.config, "cpu")
                mod.inv_freq = inv
                if hasattr(mod, "original_inv_freq"):
                    mod.original_inv_freq = inv.clone()
    m.generation_config.eos_token_id = None
    m.generation_config.pad_token_id = 0
    return m.eval()


def _reference_step(model, layer, wl, rows, n_concept=2, n_base=4):
    """The reference's own procedure on a sample: capture + mean-diff for a few prompts
    (vector_utils.py:524-545), then hook-steered HF generate for `rows` trials (model_utils.py:687-879)."""
    from oracle import ref_hooks

    cacts = ref_hooks.capture(model, wl["c_ids"][:n_concept], wl["c_mask"][:n_concept], layer)
    bacts = ref_hooks.capture(model, wl["b_ids"][:n_base], wl["b_mask"][:n_base], layer)
    vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")
    sel = list(range(rows))
    ids, mask = wl["ids"][sel], wl["mask"][sel]
    keep = int(mask.sum(1).max())  # re-pad the sample to its own width, as the tokenizer would
    ids, mask = ids[:, -keep:], mask[:, -keep:]
    out = ref_hooks.steered_generate(model, ids, mask, layer, [vecs[i % n_concept] for i in sel], STRENGTHS[-1],
                                     [wl["t0"][i] for i in sel], N_NEW)
    return out["new_ids"]


def _use_all_host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1; the CPU legs are meant to use every host core."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_baseline_port(gpu_model, layer, wl, sample_rows=2):
    threads = _use_all_host_threads()
    try:
        model = _cpu_model_from(gpu_model)
        t = time.perf_counter()
        _reference_step(model, layer, wl, sample_rows)
        dt = time.perf_counter() - t
        del model
        return {"value": sample_rows / dt, "unit": UNIT, "cores": threads, "kind": "port", "seconds": dt,
                "host_cpus": os.cpu_count(),
                "sample": f"{sample_rows} steered trials (prompt <= {T_PROMPT}, {N_NEW} greedy tokens) + extraction of "
                          f"2 concept / 4 baseline prompts, same architecture and dtype, oracle/ref_hooks.py "
                          f"(HF generate + forward hook) on the host CPU, one cold pass"}
    except Exception as exc:  # report, do not kill the GPU measurement
        return {"value": None, "unit": UNIT, "cores": threads, "kind": "port", "sample": f"failed: {exc!r}"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU baseline runs once per box
    from introspective_awareness_b200.model_utils import get_layer_at_fraction

    torch.manual_seed(0)
    _use_all_host_threads()
    model = _fast_cpu_model(args.arch)
    n_layers = model.config.num_hidden_layers
    layer = max(0, min(int(n_layers * LAYER_FRACTION), n_layers - 1))
    wl = make_workload(model.config.vocab_size, 0)
    rows = args.cpu_rows
    for _ in range(args.warmup):
        _reference_step(model, layer, wl, rows)
    t = time.perf_counter()
    for _ in range(args.steps):
        _reference_step(model, layer, wl, rows)
    dt = time.perf_counter() - t
    value = rows * args.steps / dt
    threads = torch.get_num_threads()
    sample = (f"{rows} steered trials per step (prompt <= {T_PROMPT}, {N_NEW} greedy tokens) + extraction of 2 concept / "
              f"4 baseline prompts; reference procedure restated in oracle/ref_hooks.py (HF generate + forward hook), "
              f"{args.arch} arch bf16 on the host CPU")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic (random-init weights, stub-tokenised uniform ids)",
        "config": {"workload": f"BASELINE configs[1] on the CPU, bounded sample: {sample}", "layers": [layer]},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--arch", default=ARCH)
    ap.add_argument("--cpu-rows", type=int, default=8,
                    help="steered trials per CPU step of the reference arm / cpu_baseline sample (batched like the reference batches)")
    ap.add_argument("--attention", default="auto", choices=["auto", "gqa", "hf"],
                    help="gqa: grouped-query SDPA without repeat_kv copies; hf: the model's stock attention function")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ncu-window", action="store_true", help="cudaProfilerStart at the last timed step (for ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
