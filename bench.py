"""Headline benchmark: steered trials/sec on the Llama-3.1-8B architecture (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # our arm (one process per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   # reference CPU path

A "step" is one pass of the hot path over one batch of synthetic input (BASELINE.json configs[1]):
extract 50 concept vectors against 100 baseline prompts at the steering layer (two forwards with K2's
capture tap + K1), then run 200 steered trials (50 concepts x 4 strengths, one batch, a different
(concept, alpha, t0) per row) for 32 greedy tokens each (prompt pass + 31 cached decode steps, K2 at the
layer in each).  With N > 1 every rank is an independent replica steering its own layer of the sweep
(configs[2], weak scaling): it extracts its own layer's vectors from the whole prompt sets (the reference's
batch composition, so the reference's bits at any rank count) and the token ids are gathered to rank 0 over
NCCL; `--sharded-extraction` splits the prompts over the ranks instead (`ModelWrapper.extract_vectors_sharded_ids`:
one all-reduce of the fp32 baseline sums, one all-gather of the concept rows) — faster, but each rank's forward is
then a smaller GEMM problem and the bf16 activations move with the rank count (`parity_check` reports by how much).

Prints ONE JSON line on rank 0.  `value` times the path with inputs resident in HBM; `e2e` times the
same work through the public ModelWrapper-level API with pinned HOST buffers (H2D of ids / masks /
vectors / row metadata and D2H of vectors and token ids inside the timed region).  `roofline`
is the fused inject+RMSNorm kernel's prompt-pass launch (the largest launch of ours in a step),
timed live with CUDA events on its stream.  `cpu_baseline` is the reference's CPU path (the
torch restatement in oracle/ref_hooks.py: HF generate + forward hook) on the host cores, on a
bounded sample.  Outside the timed regions the line also carries: `gpu_hook_baseline` (the reference's
own mechanism — torch-eager hooks + HF generate, one call per strength — on the same GPU and model, the
honest "before"), `n_new_100` (the same step at the reference's default of 100 new tokens),
`cold_shape_s` (what a batch shape the runtime has not seen costs), `strong_scaling` (configs[2] as
written: all layers x 200 rows dealt over the ranks), at N > 1 `parity_check` (sharded vs un-sharded
vectors, steered logits vs the hook path on every rank) and at N = 8 `other_archs` (configs[3] / [4]).
Nothing here reads /root/reference.
"""

from __future__ import annotations

import argparse
import gc
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
OTHER_ARCHS = ["gemma-2-9b", "qwen2.5-32b"]  # BASELINE.json configs[3] / configs[4]


def metric_name(arch: str) -> str:
    return METRIC if arch == ARCH else f"steered trials/sec ({arch})"


def peak_hbm():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_from_profiles(form: str):
    """dram read+write bytes per launch of the prompt-pass K2, from the committed ncu --set full summary of the
    form that was launched ("add": residual-add form, "tail": Gemma-2 block tail, "plain")."""
    for name in {"add": ["r2_k2_add_stream_4096_ncu_summary.json", "k2_add_prefill_ncu_summary.json"],
                 "tail": ["r2_k2_tail_stream_3584_ncu_summary.json"],
                 "plain": ["r2_k2_plain_tma_4096_ncu_summary.json", "k2_prefill_ncu_summary.json"]}[form]:
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", name)))
            return float(d["dram_bytes_per_launch"]), name
        except Exception:
            continue
    return None, None


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


class Arm:
    """One architecture on this rank: model, wrapper, workload, and the two step functions."""

    def __init__(self, arch, attention, rank, world, local, sharded_extraction=False):
        from introspective_awareness_b200 import parallel
        from introspective_awareness_b200.engine import SteerSpec
        from introspective_awareness_b200.model_utils import ModelWrapper, get_layer_at_fraction
        from introspective_awareness_b200.stub_tokenizer import StubTokenizer
        from introspective_awareness_b200.synthetic import build_model

        self.arch, self.rank, self.world = arch, rank, world
        self.sharded_extraction = bool(sharded_extraction) and world > 1
        self.parallel, self.SteerSpec = parallel, SteerSpec
        dev = self.dev = torch.device("cuda", local)
        torch.manual_seed(0)
        self.model = build_model(arch, dtype=torch.bfloat16, device=str(dev), seed=None)
        self.w = ModelWrapper(arch, device=str(dev), dtype=torch.bfloat16, model=self.model,
                              tokenizer=StubTokenizer(vocab_size=self.model.config.vocab_size), attention=attention)
        eng = self.eng = self.w.engine
        n_layers = eng.n_layers
        base_layer = get_layer_at_fraction(self.w, LAYER_FRACTION)
        # rank r steers its own layer of the sweep (weak scaling; N=1 -> the single layer of configs[1])
        self.sweep = [(base_layer + r) % n_layers for r in range(world)] if world > 1 else [base_layer]
        self.my_layer = self.sweep[rank]
        wl = self.wl = make_workload(self.model.config.vocab_size, rank)
        self.rows = wl["rows"]
        on = lambda t: t.to(dev)
        self.c_ids, self.c_mask, self.b_ids, self.b_mask = on(wl["c_ids"]), on(wl["c_mask"]), on(wl["b_ids"]), on(wl["b_mask"])
        self.ids_d, self.mask_d = on(wl["ids"]), on(wl["mask"])
        pads = T_PROMPT - wl["lens"]
        self.t0_abs = (torch.tensor(wl["t0"]) + pads).to(torch.int32)
        self.row_vec_d = torch.tensor(wl["row_vec"], dtype=torch.int32, device=dev)
        self.row_alpha_d = torch.tensor(wl["row_alpha"], dtype=torch.float32, device=dev)
        self.t0_d = self.t0_abs.to(dev)
        pin = lambda t: t.clone().pin_memory()  # pinned host copies for the end-to-end arm
        self.h_c_ids, self.h_c_mask, self.h_b_ids, self.h_b_mask = pin(wl["c_ids"]), pin(wl["c_mask"]), pin(wl["b_ids"]), pin(wl["b_mask"])
        self.h_ids, self.h_mask = pin(wl["ids"]), pin(wl["mask"])

    def session(self, n_new=N_NEW):
        return self.eng.session(self.rows, T_PROMPT, n_new, padded=True, n_vec_cap=64, vec_dtype=torch.bfloat16)

    def extract(self, layers, host=False, sharded=None):
        """-> [len(layers), 50, H] device vectors (inputs: the HBM-resident copies, or the pinned host ones).
        Default: every rank extracts the layers it needs from the WHOLE prompt sets — the reference's batch
        composition, hence the reference's bits whatever the rank count.  sharded (--sharded-extraction): the prompts
        split over the ranks through the product API (a collective call; see parity_check for what that costs in bits)."""
        ci, cm, bi, bm = ((self.h_c_ids, self.h_c_mask, self.h_b_ids, self.h_b_mask) if host
                          else (self.c_ids, self.c_mask, self.b_ids, self.b_mask))
        if self.sharded_extraction if sharded is None else sharded:
            _, vr, _ = self.w.extract_vectors_sharded_ids(ci, cm, bi, bm, layers)
        else:
            _, vr, _ = self.eng.extract_vectors(ci, cm, bi, bm, layers, want_f32=False, want_norm=False)
        return vr

    def step_layers(self):
        """Layers one step extracts for: this rank's own, or (prompt-sharded: a collective) every rank's."""
        return sorted(set(self.sweep)) if self.sharded_extraction else [self.my_layer]

    def step_resident(self, n_new=N_NEW, layer=None, table=None):
        if table is None:
            uniq = self.step_layers()
            table = self.extract(uniq)[uniq.index(self.my_layer)]
        steer = self.SteerSpec(self.my_layer if layer is None else layer, table, self.row_vec_d, self.row_alpha_d, self.t0_d)
        out = self.session(n_new).generate(self.ids_d, self.mask_d, steer, temperature=0.0, eos_token_id=None, pad_token_id=0)
        if self.world > 1:
            self.parallel.gather_to_rank0(out)
        return out

    def step_e2e(self):
        """Public API with host buffers: ids/masks from pinned host memory, vectors returned to the host
        (as the reference's extraction does), generation fed with host vectors, ids read back."""
        uniq = self.step_layers()
        vec_host = self.extract(uniq, host=True)[uniq.index(self.my_layer)].cpu()
        wl = self.wl
        out = self.w.steered_generate_ids(self.h_ids, self.h_mask, self.my_layer, vec_host, wl["row_vec"], wl["row_alpha"], wl["t0"],
                                          N_NEW, temperature=0.0)
        host_ids = out.cpu()
        if self.world > 1:
            self.parallel.gather_to_rank0(out)
        return host_ids

    def barrier(self):
        if self.world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps):
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
        return float(ms.item())

    def free(self):
        self.eng._session = None
        self.eng._extract_graphs.clear()
        for name in ("w", "eng", "model"):
            delattr(self, name)
        gc.collect()
        torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------
# checks and context figures (all outside the timed regions)
# ---------------------------------------------------------------------------------------------

def parity_check(arm: Arm) -> dict:
    """N > 1, before timing.  (a) this rank's vectors (extracted from the whole prompt sets, as every step does by
    default) against the hook path's capture + CPU arithmetic on the same GPU; (b) this rank's steered rows against the
    hook path on a 16-row subset of its own batch: first-step logits and the greedy tokens; (c) for the record, what the
    optional prompt-sharded extraction (K1a -> all-reduce -> all-gather -> K1b) does to the bits: each rank then
    forwards a smaller batch, cuBLAS picks other tiles for the smaller M, and the bf16 activations move."""
    from oracle import ref_hooks

    uniq = sorted(set(arm.sweep))
    sharded = arm.extract(uniq, sharded=True)                     # collective
    _, whole, _ = arm.eng.extract_vectors(arm.c_ids, arm.c_mask, arm.b_ids, arm.b_mask, uniq, want_f32=False, want_norm=False)
    a, b = sharded.float(), whole.float()
    rel = ((a - b).norm(dim=-1) / b.norm(dim=-1)).max().item()
    differing = (sharded != whole).float().mean().item()
    table = whole[uniq.index(arm.my_layer)].clone()
    hook_vecs = ref_hooks.concept_vectors(ref_hooks.capture(arm.model, arm.wl["c_ids"], arm.wl["c_mask"], arm.my_layer),
                                          ref_hooks.capture(arm.model, arm.wl["b_ids"], arm.wl["b_mask"], arm.my_layer), "baseline")
    vec_rel = ((table.float().cpu() - hook_vecs.float()).norm(dim=-1) / hook_vecs.float().norm(dim=-1)).max().item()
    wl = arm.wl
    out = arm.w.steered_generate_ids(arm.ids_d, arm.mask_d, arm.my_layer, table, wl["row_vec"], wl["row_alpha"], wl["t0"], N_NEW,
                                     temperature=0.0, keep_logits=True).clone()
    ours_logits = torch.stack(arm.w.last_session.step_logits)
    sub = list(range(16))                                          # rows 0..15 share strength STRENGTHS[0]
    ref = ref_hooks.steered_generate(arm.model, wl["ids"][sub], wl["mask"][sub], arm.my_layer, [table[wl["row_vec"][i]].cpu() for i in sub],
                                     STRENGTHS[0], [wl["t0"][i] for i in sub], N_NEW, want_logits=True)
    first = (ours_logits[0, :16].cpu() - ref["logits"][0]).abs().max().item()
    same = (out[:16].cpu() == ref["new_ids"]).all(dim=1).float().mean().item()
    arm.w.last_session.logits_buf = None  # the timed steps run the graphs without the logits copy
    arm.w.last_session.graphs.clear()
    stats = torch.tensor([rel, differing, first, 1.0 - same, vec_rel], device=arm.dev)
    if arm.world > 1:
        torch.distributed.all_reduce(stats, op=torch.distributed.ReduceOp.MAX)
    rel, differing, first, miss, vec_rel = stats.tolist()
    return {"vectors_vs_hook_path_rel_l2_max": vec_rel, "first_step_logits_vs_hook_path_max_abs": first,
            "rows_with_identical_32_token_paths": 1.0 - miss, "rows_checked_per_rank": 16, "ranks": arm.world,
            "ok": bool(vec_rel <= 1e-3 and first <= 2e-2),
            "extraction_in_timed_steps": "prompt-sharded" if arm.sharded_extraction else "whole prompt sets on every rank",
            "prompt_sharded_extraction": {"vs_unsharded_vectors_rel_l2_max": rel, "fraction_of_elements_differing": differing,
                                          "within_north_star_1e-3": bool(rel <= 1e-3)},
            "note": "max over ranks; vectors: each rank's extraction against ref_hooks capture + CPU arithmetic on the same GPU "
                    "(north_star: <= 1e-3); logits: 200-row batch of ours against a 16-row hook-path batch of the same rows "
                    "(<= 2e-2); prompt_sharded_extraction: the optional sharded protocol against the un-sharded result — a "
                    "rank-count-dependent difference, because each rank's forward is a smaller GEMM problem"}


def gpu_hook_baseline(arm: Arm, reps: int = 2) -> dict:
    """The reference's own mechanism on this GPU and model: hook capture with a full-depth forward per prompt set
    and `.cpu()` (model_utils.py:293-345), CPU vector arithmetic (vector_utils.py:541-548), then hook-steered HF
    `generate` — one call per strength because `strength` is a scalar there (model_utils.py:692, :737), i.e. four
    calls of 50 rows for the 200-row step."""
    from oracle import real_ref, ref_hooks

    wl, model, layer = arm.wl, arm.model, arm.my_layer
    kind = reference_kind()

    def step():
        if kind == "reference":  # the reference's own ModelWrapper and vector_utils, on this GPU
            w = _real_wrapper(model)
            vecs = real_ref.concept_vectors(w, wl["c_ids"], wl["c_mask"], wl["b_ids"], wl["b_mask"], layer)
        else:
            cacts = ref_hooks.capture(model, wl["c_ids"], wl["c_mask"], layer)
            bacts = ref_hooks.capture(model, wl["b_ids"], wl["b_mask"], layer)
            vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")
        for k, a in enumerate(STRENGTHS):
            sel = list(range(k * N_CONCEPTS, (k + 1) * N_CONCEPTS))
            ids, mask = wl["ids"][sel], wl["mask"][sel]
            if kind == "reference":
                real_ref.steered_generate(w, ids, mask, layer, [vecs[i % N_CONCEPTS] for i in sel], a, [wl["t0"][i] for i in sel], N_NEW)
                continue
            keep = int(mask.sum(1).max())  # each call is padded to its own width, as the tokenizer would
            ref_hooks.steered_generate(model, ids[:, -keep:], mask[:, -keep:], layer, [vecs[i % N_CONCEPTS] for i in sel], a,
                                       [wl["t0"][i] for i in sel], N_NEW)

    step()
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(reps):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / reps
    return {"value": arm.rows / dt, "unit": UNIT, "ms_per_step": dt * 1e3, "calls_per_step": len(STRENGTHS), "steps_timed": reps, "kind": kind,
            "what": ("the reference's own ModelWrapper.generate_batch_with_multi_steering and vector_utils.extract_concept_vectors_batch "
                     "(oracle/_ref)" if kind == "reference" else "oracle/ref_hooks.py (restated reference)")
                    + " on the same GPU and model: torch-eager forward hooks + HF generate (DynamicCache, no CUDA graphs, stock "
                      "attention), extraction by full-depth hooked forwards + CPU arithmetic; same 200 rows, 32 greedy tokens, "
                      "50 + 100 extraction prompts"}


def cold_shape(arm: Arm) -> dict:
    """What a batch shape the runtime has not seen costs: the first `steered_generate_ids` at a new padded width
    captures max_new_tokens - 1 step graphs; the second call of the same shape is the steady state.  Kept sessions
    (DecoderEngine.max_sessions) and layer-independent graphs bound how often this is paid in a sweep."""
    out = {}
    wl = arm.wl
    table = arm.extract([arm.my_layer])[0].clone()
    for n_new, width in ((N_NEW, T_PROMPT - 1), (100, T_PROMPT - 2)):
        ids, mask = wl["ids"][:, -width:].contiguous(), wl["mask"][:, -width:].contiguous()
        t0 = [max(0, t - (T_PROMPT - width)) for t in wl["t0"]]
        times = []
        for _ in range(2):
            torch.cuda.synchronize()
            t = time.perf_counter()
            arm.w.steered_generate_ids(ids, mask, arm.my_layer, table, wl["row_vec"], wl["row_alpha"], t0, n_new, temperature=0.0)
            torch.cuda.synchronize()
            times.append(time.perf_counter() - t)
        t = time.perf_counter()  # another LAYER on the warm shape: no new graphs
        arm.w.steered_generate_ids(ids, mask, (arm.my_layer + 1) % arm.eng.n_layers, table, wl["row_vec"], wl["row_alpha"], t0, n_new,
                                   temperature=0.0)
        torch.cuda.synchronize()
        out[f"n_new_{n_new}"] = {"first_call_s": times[0], "second_call_s": times[1], "other_layer_same_shape_s": time.perf_counter() - t,
                                 "graphs_captured": n_new - 1}
    out["sessions_kept"] = arm.eng.max_sessions
    return out


def strong_scaling(arm: Arm) -> dict:
    """configs[2] as written: ALL layers x 200 rows, the layer axis dealt round-robin over the ranks, one
    extraction for every layer (prompt-sharded at N > 1); seconds for the whole sweep, max over ranks."""
    n_layers = arm.eng.n_layers
    mine = [l for l in range(n_layers) if l % arm.world == arm.rank]

    def sweep():
        if arm.sharded_extraction:
            vecs = dict(enumerate(arm.extract(list(range(n_layers)))))
        else:  # every rank: one extraction for its own layers, from the whole prompt sets
            vecs = dict(zip(mine, arm.extract(mine)))
        for l in mine:
            arm.step_resident(layer=l, table=vecs[l])

    sweep()  # warm: extraction graph of the all-layer capture
    ms = arm.timed(sweep, 1)
    total = n_layers * arm.rows
    return {"layers": n_layers, "rows_total": total, "layers_per_rank": len(mine), "seconds": ms / 1e3, "value": total / (ms / 1e3),
            "unit": UNIT, "note": "fixed total work (BASELINE configs[2]); `value` on the headline line is the weak-scaling figure"}


def measure_other_arch(arch, args, rank, world, local) -> dict:
    arm = Arm(arch, args.attention, rank, world, local, args.sharded_extraction)
    try:
        for _ in range(3):
            arm.step_resident()
        arm.eng.k2_events = []
        ms = arm.timed(arm.step_resident, 3)
        k2_ms = [a.elapsed_time(b) for a, b, _ in arm.eng.k2_events]
        arm.eng.k2_events = None
        return {"metric": metric_name(arch), "value": arm.rows * world * 3 / (ms / 1e3), "unit": UNIT, "ms_per_step": ms / 3, "steps": 3,
                "warmup": 3, "layers": arm.sweep, "rows_per_gpu_per_step": arm.rows, "n_gpus": world,
                "k2_prompt_pass_us": (sum(k2_ms) / len(k2_ms) * 1e3) if k2_ms else None,
                "engine_block": "native" if (arm.eng.native_block or arm.eng.gemma_native_attention) else "hf",
                "attention": arm.eng.attention}
    finally:
        arm.free()


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------

def run_ours(args):
    from introspective_awareness_b200 import parallel

    rank, world, local = parallel.init_from_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (our arm) needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    arch = args.arch
    arm = Arm(arch, args.attention, rank, world, local, args.sharded_extraction)
    eng, rows = arm.eng, arm.rows

    parity = None
    if world > 1:
        arm.step_resident()  # first call: builds the session and the step graphs
        parity = parity_check(arm)

    step_resident = arm.step_resident
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

    ms_res = arm.timed(step_resident, args.steps)
    launches_after = dict(eng.counters)
    if args.ncu_window:
        torch.cuda.profiler.stop()
    k2_events = eng.k2_events
    eng.k2_events = None
    clocks = sampler.stop() if sampler else None
    for _ in range(2):
        arm.step_e2e()
    ms_e2e = arm.timed(arm.step_e2e, args.steps)

    trials = rows * world * args.steps
    value = trials / (ms_res / 1e3)
    e2e_value = trials / (ms_e2e / 1e3)

    # roofline of the dominant launch of ours: K2 over the prompt pass [rows, T_PROMPT, H]
    k2_ms = [a.elapsed_time(b) for a, b, _ in k2_events]
    H = eng.hidden
    tail = eng.fuse_gemma_tail  # Gemma-2: the block-tail variant reads mlp_out as a fifth stream and a second weight
    # Llama / Qwen2 (engine-driven block): the launch is the residual-add form — the MLP output is a fifth stream
    fifth = tail or eng.native_block
    form = "tail" if tail else ("add" if eng.native_block else "plain")
    nvec = rows * T_PROMPT
    algo = nvec * H * 2 * (5 if fifth else 4) + H * 2 * (2 if tail else 1)   # the table row counted as a stream
    dram_needed = nvec * H * 2 * (4 if fifth else 3)                            # the <= 0.5 MB table stays in L2
    peak, peak_src = peak_hbm()
    k2_avg_ms = sum(k2_ms) / max(len(k2_ms), 1)
    sec = k2_avg_ms / 1e3
    # only positions >= t0 are steered in this launch: they alone read a table row (L2) — and in the plain form write h'
    steered_vecs = int((T_PROMPT - arm.t0_abs.clamp(max=T_PROMPT)).sum())
    moved = (nvec * (4 if fifth else 2) + (0 if fifth else steered_vecs)) * H * 2
    traffic, traffic_src = traffic_from_profiles(form)
    # which implementation the library picks for this launch (csrc/ia_inject_rmsnorm.cu, k2_entry)
    impl_name = ("ia::k2_stream_kernel: row in registers, quarter-row cp.async.bulk landing slots"
                 if H in (3584, 4096, 5120) and (form != "plain" or H == 5120) else "ia::k2_tma_kernel: cp.async.bulk staged rows")

    h2d = sum(t.numel() * t.element_size() for t in (arm.h_c_ids, arm.h_c_mask, arm.h_b_ids, arm.h_b_mask, arm.h_ids, arm.h_mask)) \
        + N_CONCEPTS * H * 2 + rows * (4 + 4 + 4)
    d2h = N_CONCEPTS * H * 2 + rows * N_NEW * 8
    # our kernels launched inside the value-timed region on this rank (counted by the engine: eager
    # launches directly, graph replays by the launches recorded at capture)
    launches_by_kernel = {k: launches_after[k] - launches_before[k] for k in launches_before}
    launches = sum(launches_by_kernel.values())

    # ---- context figures, outside the timed regions ----
    extras = {}
    if not args.no_extras:
        if world == 1:
            # the same step at the reference's default of 100 new tokens (detect_injected_thoughts.py:73)
            for _ in range(2):
                arm.step_resident(n_new=100)
            ms100 = arm.timed(lambda: arm.step_resident(n_new=100), 2)
            extras["n_new_100"] = {"value": rows * 2 / (ms100 / 1e3), "unit": UNIT, "ms_per_step": ms100 / 2, "new_tokens": 100, "steps": 2,
                                   "warmup": 2}
            extras["cold_shape_s"] = cold_shape(arm)
        extras["strong_scaling"] = strong_scaling(arm)
        if world == 1:
            eng._session = None  # the hook path's DynamicCache and activations need the room
            hook = gpu_hook_baseline(arm)
            hook["ours_over_hook_path"] = value / hook["value"]
            extras["gpu_hook_baseline"] = hook

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:  # reported at N=1 only
        cpu = cpu_baseline_port(arm.model, arm.my_layer, arm.wl, sample_rows=args.cpu_rows)

    sweep, fused, attention, native = arm.sweep, dict(eng._fused_checked), eng.attention, eng.native_block
    block_kind = ("gemma-2: engine attention + table MLP + block-tail K2" if getattr(eng, "gemma_native_attention", False) and eng.table_mlp
                  else "hf")
    arm_sharded = arm.sharded_extraction
    if (world == 8 or args.other_archs) and arch == ARCH and not args.no_extras:
        del eng
        arm.free()
        extras["other_archs"] = {a: measure_other_arch(a, args, rank, world, local) for a in OTHER_ARCHS}

    if rank == 0:
        line = {
            "metric": metric_name(arch), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_res / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (random-init weights, stub-tokenised uniform ids)",
            "config": {
                "workload": f"BASELINE configs[1]: {arch} arch, {N_CONCEPTS} concepts x {len(STRENGTHS)} strengths at one "
                            f"layer, one {rows}-row steered batch per GPU, prompt {T_PROMPT} (real {T_MIN}..{T_PROMPT}, "
                            f"left-padded), {N_NEW} greedy tokens, extraction {N_CONCEPTS}+{N_BASE} prompts per step",
                "layers": sweep, "rows_per_gpu_per_step": rows, "new_tokens": N_NEW, "attention": attention,
                "engine_block": "native" if native else block_kind,
                "single_gemm_projections": {f"{k}{list(shape)}": ok for (k, shape), ok in sorted(fused.items())},
                "parallelism": ("independent replicas, layer axis round-robin; extraction "
                                + ("prompt-sharded (NCCL all-reduce / all-gather)" if arm_sharded else "from the whole prompt sets on every rank")
                                if world > 1 else "single replica"),
                "l2": "inputs larger than L2: each step streams 16 GB of weights per decode step and ~1 GB of KV/activations",
            },
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "gpu_launches_by_kernel": launches_by_kernel,
            "roofline": {"kernel": {"add": "ia_add_inject_rmsnorm (residual add + inject + RMSNorm; " + impl_name + "), prompt-pass launch",
                                    "tail": "ia_gemma2_tail_inject_rmsnorm (post-FFN norm + residual add + inject + next RMSNorm; "
                                            + impl_name + "), prompt-pass launch",
                                    "plain": "ia_inject_rmsnorm (fused inject+RMSNorm; " + impl_name + "), prompt-pass launch"}[form],
                         "bound": "hbm",
                         # frac on the bytes DRAM has to move (3 streams plain, 4 with a branch row): the table row is an L2 hit
                         "achieved": dram_needed / sec / 1e9 if k2_ms else None, "peak": peak, "unit": "GB/s",
                         "frac": dram_needed / sec / 1e9 / peak if k2_ms else None,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "dram_necessary_bytes_per_launch": dram_needed,
                         "survey_8d_bytes_per_launch": algo, "survey_8d_achieved": algo / sec / 1e9 if k2_ms else None,
                         "survey_8d_frac": algo / sec / 1e9 / peak if k2_ms else None,
                         "us_per_launch": k2_avg_ms * 1e3,
                         "steered_fraction_of_vectors": steered_vecs / nvec,
                         "bytes_moved_per_launch": moved,
                         "launches_timed": len(k2_ms),
                         "share_of_step": (k2_avg_ms * len(k2_ms) / ms_res) if k2_ms else None,
                         "streams_per_vector": 5 if fifth else 4,
                         "note": "achieved / frac use the bytes DRAM must move per launch: residual in, h' out, y out for every "
                                 "vector, plus the branch row in when the layer's residual add (or Gemma-2's block tail) is folded "
                                 "into the launch; the concept-table row (<= 0.5 MB table) is an L2 hit and is left out. The "
                                 "survey_8d_* figures count it as a stream, as SURVEY.md 8(d)'s formula does. In a bench step the "
                                 "branch row was just written by the preceding GEMM and is partly L2-resident, so this live figure "
                                 "can exceed the cold ladder's (tools/kbench.py --r2 -> profiles/r2_kbench_ladder.jsonl); `traffic` "
                                 "is the DRAM byte count per launch of the ncu --set full capture named in traffic_source"},
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        if parity is not None:
            line["parity_check"] = parity
        line.update(extras)
        emit(line)
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


SAME_CONFIG_NOTE = (
    "not the same amount of work per step as the GPU arm: the CPU arm extracts from the real 50 + 100 prompts but decodes a "
    "sample of {rows} of the 200 steered trials per step (one 200-row CPU step is ~400 TFLOP of prompt pass alone: hours); "
    "trials/s is per trial on both arms, but a CPU decode step streams the same 16 GB of weights for {rows} rows as it would "
    "for 200, so a full-batch CPU step would show more trials/s than this sample does, and a ratio against it is a "
    "GPU-vs-CPU-sample figure. The like-for-like 'before' is gpu_hook_baseline: the reference's hook path on the same GPU.")


_REF_WRAPPERS = {}


def reference_kind() -> str:
    """"reference": oracle/_ref holds the real reference (compiled by oracle/build_ref.py) and the legs below run ITS
    ModelWrapper / vector_utils; "port": they run the restatement in oracle/ref_hooks.py."""
    from oracle import real_ref

    return "reference" if real_ref.load() is not None else "port"


def _real_wrapper(model):
    from oracle import real_ref

    w = _REF_WRAPPERS.get(id(model))
    if w is None:
        p = next(model.parameters())
        w = _REF_WRAPPERS[id(model)] = real_ref.wrap(model, str(p.device), p.dtype)
    return w


def _reference_step(model, layer, wl, rows):
    """The reference's own procedure: capture + mean-diff for the full 50 concept and 100 baseline prompts
    (vector_utils.py:524-545; two full-depth hooked forwards), then hook-steered HF generate for a sample of
    `rows` trials (model_utils.py:687-879) — through the real reference's own functions when oracle/_ref is there."""
    from oracle import real_ref, ref_hooks

    sel = list(range(rows))
    if real_ref.load() is not None:
        w = _real_wrapper(model)
        vecs = real_ref.concept_vectors(w, wl["c_ids"], wl["c_mask"], wl["b_ids"], wl["b_mask"], layer)
        return real_ref.steered_generate(w, wl["ids"][sel], wl["mask"][sel], layer, [vecs[wl["row_vec"][i]] for i in sel],
                                         STRENGTHS[0], [wl["t0"][i] for i in sel], N_NEW)
    cacts = ref_hooks.capture(model, wl["c_ids"], wl["c_mask"], layer)
    bacts = ref_hooks.capture(model, wl["b_ids"], wl["b_mask"], layer)
    vecs = ref_hooks.concept_vectors(cacts, bacts, "baseline")
    ids, mask = wl["ids"][sel], wl["mask"][sel]
    keep = int(mask.sum(1).max())  # re-pad the sample to its own width, as the tokenizer would
    ids, mask = ids[:, -keep:], mask[:, -keep:]
    out = ref_hooks.steered_generate(model, ids, mask, layer, [vecs[wl["row_vec"][i]] for i in sel], STRENGTHS[0],
                                     [wl["t0"][i] for i in sel], N_NEW)
    return out["new_ids"]


def _sample_text(rows, arch):
    how = ("the reference's own vector_utils.extract_concept_vectors_batch and ModelWrapper.generate_batch_with_multi_steering "
           "(oracle/_ref, compiled from the reference by oracle/build_ref.py; prompts = the synthetic ids written out as numbers)"
           if reference_kind() == "reference" else "reference procedure restated in oracle/ref_hooks.py (HF generate + forward hook)")
    return (f"{rows} steered trials per step (prompt <= {T_PROMPT}, {N_NEW} greedy tokens) + extraction of the full {N_CONCEPTS} concept / "
            f"{N_BASE} baseline prompts; {how}, {arch} arch bf16 on the host CPU")


def _use_all_host_threads() -> int:
    """torchrun exports OMP_NUM_THREADS=1; the CPU legs are meant to use every host core."""
    n = os.cpu_count() or 1
    try:
        n = len(os.sched_getaffinity(0))
    except Exception:
        pass
    torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_baseline_port(gpu_model, layer, wl, sample_rows=8):
    threads = _use_all_host_threads()
    try:
        model = _cpu_model_from(gpu_model)
        t = time.perf_counter()
        _reference_step(model, layer, wl, sample_rows)
        dt = time.perf_counter() - t
        arch = getattr(gpu_model.config, "model_type", "?")
        del model
        return {"value": sample_rows / dt, "unit": UNIT, "cores": threads, "kind": reference_kind(), "seconds": dt,
                "host_cpus": os.cpu_count(), "sample": _sample_text(sample_rows, arch) + ", one cold pass",
                "same_config_note": SAME_CONFIG_NOTE.format(rows=sample_rows)}
    except Exception as exc:  # report, do not kill the GPU measurement
        return {"value": None, "unit": UNIT, "cores": threads, "kind": reference_kind(), "sample": f"failed: {exc!r}"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU baseline runs once per box
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
    sample = _sample_text(rows, args.arch)
    emit({
        "impl": "reference", "metric": metric_name(args.arch), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic (random-init weights, stub-tokenised uniform ids)",
        "config": {"workload": f"BASELINE configs[1] on the CPU, bounded sample: {sample}", "layers": [layer],
                   "same_config_note": SAME_CONFIG_NOTE.format(rows=rows)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": reference_kind(), "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    })


_REAL_STDOUT = None


def _stdout_for_the_json_line_only():
    """stdout carries the one JSON line and nothing else: whatever libraries print there meanwhile (NCCL's version banner
    when the box sets NCCL_DEBUG=VERSION — a plain printf, NCCL_DEBUG_FILE does not move it) is sent to stderr by pointing
    file descriptor 1 at stderr until `emit` puts it back."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    global _REAL_STDOUT
    sys.stdout.flush()
    if _REAL_STDOUT is not None:
        os.dup2(_REAL_STDOUT, 1)
        os.close(_REAL_STDOUT)
        _REAL_STDOUT = None
    print(json.dumps(line), flush=True)


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
    ap.add_argument("--no-extras", action="store_true", help="skip the context figures measured outside the timed regions")
    ap.add_argument("--other-archs", action="store_true", help="also run one short step each of configs[3] / [4] (default at --gpus 8)")
    ap.add_argument("--sharded-extraction", action="store_true",
                    help="N > 1: split the extraction prompts over the ranks (NCCL all-reduce of fp32 sums, all-gather of concept "
                         "rows) instead of extracting from the whole prompt sets on every rank; faster, not rank-count-invariant in bits")
    ap.add_argument("--ncu-window", action="store_true", help="cudaProfilerStart at the last timed step (for ncu)")
    args = ap.parse_args()
    _stdout_for_the_json_line_only()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
