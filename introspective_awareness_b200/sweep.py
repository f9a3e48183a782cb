"""Layer x strength sweep on the batched runtime, writing the reference's on-disk formats
 so that the reference's `--reevaluate-judge`, plotting and transcript tools
(detect_injected_thoughts.py:1400-1506, :560, :771, :1080) can consume the output unchanged:

    <out>/<model>/layer_{fraction:.2f}_strength_{strength}/results.json   {"results", "metrics", "n_samples"}
    <out>/<model>/layer_{fraction:.2f}_strength_{strength}/results.csv    one row per trial
    <out>/<model>/vectors/layer_{fraction:.2f}/{concept}.pt               bare torch.save of the vector

(reference detect_injected_thoughts.py:1651, :1780-1787, :2136-2157; eval_utils.py:894-919).

What differs from the reference's sweep (detect_injected_thoughts.py:1551-1561, :1645-1647):
all layer fractions are extracted from ONE pair of forwards (K1 over all layers), each layer's
concept x strength x trial grid is ONE steered batch sequence with per-row strength (K2), control
trials are generated once per layer instead of once per (layer, strength), and layers are dealt
round-robin to the ranks of a `torch.distributed` job (left-over layers shared batch by batch).  The LLM judge stays outside (host-side,
untimed): metrics here are the reference's keyword metrics (`check_concept_mentioned`).
"""

from __future__ import annotations

import csv
import json
from pathlib import Path
from typing import Dict, List, Optional, Sequence

import torch

from . import grid, parallel
from . import vector_utils as vu
from .model_utils import get_layer_at_fraction

RESULT_FIELDS = ["concept", "trial", "response", "injected", "layer", "layer_fraction", "strength", "detected", "trial_type"]


def config_dir(base: Path, layer_fraction: float, strength: float) -> Path:
    return Path(base) / f"layer_{layer_fraction:.2f}_strength_{strength}"


def keyword_metrics(results: Sequence[dict]) -> Dict[str, float]:
    """What the reference driver writes as a configuration's metrics when its LLM judge is off (the judge stays
    outside this package): keyword rates from the `detected` flag, and zeros for the three figures only the judge
    can produce (detect_injected_thoughts.py:2108-2120, names and order as in :2139-2150)."""
    def rate(rows):
        rows = list(rows)
        return sum(1 for r in rows if r["detected"]) / len(rows) if rows else 0

    inj = [r for r in results if r["injected"] and r["trial_type"] == "injection"]
    ctl = [r for r in results if not r["injected"] and r["trial_type"] == "control"]
    frc = [r for r in results if r["trial_type"] == "forced_injection"]
    return {"detection_hit_rate": rate(inj), "detection_false_alarm_rate": rate(ctl), "detection_accuracy": 0,
            "identification_accuracy_given_claim": 0, "combined_detection_and_identification_rate": 0,
            "forced_identification_accuracy": rate(frc)}


def write_config(base: Path, layer_fraction: float, layer_idx: int, strength: float, results: List[dict],
                 temperature: float, max_tokens: int, extra_metrics: Optional[dict] = None) -> Path:
    """results.json + results.csv for one (layer fraction, strength) configuration."""
    d = config_dir(base, layer_fraction, strength)
    d.mkdir(parents=True, exist_ok=True)
    metrics = dict(keyword_metrics(results), layer_fraction=layer_fraction, layer_idx=layer_idx, strength=strength,
                   temperature=temperature, max_tokens=max_tokens)
    metrics.update(extra_metrics or {})
    with open(d / "results.json", "w") as f:
        json.dump({"results": results, "metrics": metrics, "n_samples": len(results)}, f, indent=2)
    with open(d / "results.csv", "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=RESULT_FIELDS, extrasaction="ignore")
        w.writeheader()
        w.writerows(results)
    return d


def save_vectors(base: Path, layer_fraction: float, vectors: Dict[str, torch.Tensor]) -> None:
    d = Path(base) / "vectors" / f"layer_{layer_fraction:.2f}"
    d.mkdir(parents=True, exist_ok=True)
    for concept, vec in vectors.items():
        torch.save(vec, d / f"{concept}.pt")


def run_sweep(model, concepts: Sequence[str], baseline_words: Sequence[str], layer_fractions: Sequence[float],
              strengths: Sequence[float], n_trials: int = 30, max_new_tokens: int = 100, temperature: float = 1.0,
              batch_size: int = 256, output_dir: Optional[str] = None, extraction_method: str = "baseline",
              with_forced: bool = True, save_vector_files: bool = True, sharded_extraction: bool = False
              ) -> Dict[tuple, List[dict]]:
    """Runs the whole grid and (if output_dir) writes the reference's files.  Trial split as the reference:
    n_injection = n_trials // 2, n_control = n_trials - n_injection, forced = n_injection, numbered 1..n_injection,
    then n_injection + t, then n_injection + n_control + t (detect_injected_thoughts.py:1792-1799, :1917, :1984).

    In a `torch.distributed` job (one process per GPU, `parallel.init_from_env`) this is a collective call: each
    rank steers the layer fractions dealt to it round-robin — the same number for every rank — and writes their
    files; the fractions left over (n % world; all of them when there are fewer layers than ranks) are shared, their
    row batches dealt over all ranks and put together and written by rank 0 (42 Gemma-2 layers on 8 GPUs: 5 whole
    layers per rank + 2 shared, not 6 / 5).  The result rows are gathered — rank 0 returns every configuration, the
    other ranks those of their whole layers.  Extraction: by default every rank
    extracts the layers IT steers from the whole prompt sets (the reference's batch composition, hence the
    reference's vectors whatever the rank count; rank 0 also extracts what it needs for the vector files);
    sharded_extraction=True splits the prompts over the ranks instead (fp32 sums all-reduced, concept rows
    all-gathered) — faster, but every rank's forward is then a smaller GEMM problem for which cuBLAS may pick other
    tiles, so the bf16 activations, and with them the vectors, depend on the rank count (measured at 8 ranks on
    Llama-3.1-8B dimensions: 4e-2 relative L2 against the un-sharded vectors; identical at 2 ranks)."""
    rank, world = parallel.world()
    concepts = list(concepts)
    fractions = list(layer_fractions)
    layer_of = {f: get_layer_at_fraction(model, f) for f in fractions}
    uniq_layers = sorted(set(layer_of.values()))
    whole, shared = parallel.shard_whole_and_shared(fractions, world)
    mine = whole[rank]
    if sharded_extraction and world > 1:
        need = uniq_layers
    else:
        keep_all = rank == 0 and output_dir is not None and save_vector_files
        need = uniq_layers if keep_all else sorted({layer_of[f] for f in mine + shared})
    by_layer = vu.extract_concept_vectors_multi_layer(model, concepts, list(baseline_words), need,
                                                      extraction_method=extraction_method,
                                                      sharded=sharded_extraction and world > 1) if need else {}
    base = None
    if output_dir is not None:
        base = Path(output_dir) / str(model.model_name).replace("/", "_")
        if save_vector_files and rank == 0:
            for f in fractions:
                save_vectors(base, f, by_layer[layer_of[f]])
    n_inj = n_trials // 2
    n_ctl = n_trials - n_inj
    rows = grid.layer_trials(concepts, strengths, n_inj, n_ctl, n_inj if with_forced else 0, dedupe_control=True)

    def configurations(f, results) -> Dict[tuple, List[dict]]:
        """One entry (and one pair of files) per (layer fraction, strength), rows in the driver's order: injection,
        control, forced (:1797-1799); the strength-independent control trials are stored in every strength's file."""
        control = [r for r in results if r["trial_type"] == "control"]
        cfgs = {}
        for a in strengths:
            steered = [r for r in results if r["trial_type"] != "control" and r["strength"] == float(a)]
            cfg = [r for r in steered if r["trial_type"] == "injection"]
            cfg += [dict(r, strength=float(a)) for r in control]
            cfg += [r for r in steered if r["trial_type"] == "forced_injection"]
            cfgs[(f, float(a))] = cfg
            if base is not None:
                write_config(base, f, layer_of[f], float(a), cfg, temperature, max_new_tokens)
        return cfgs

    out: Dict[tuple, List[dict]] = {}
    for f in mine:
        out.update(configurations(f, grid.run_layer(model, layer_of[f], by_layer[layer_of[f]], rows, max_new_tokens,
                                                    temperature, batch_size, layer_fraction=f)))
    # the layers the round-robin deal leaves over: their row batches are dealt over ALL ranks (batch u of the list of
    # (layer, batch) units to rank u % world), gathered, and put together — and written — by rank 0
    n_batches = len(list(grid.chunk(rows, batch_size)))
    units = [(f, b) for f in shared for b in range(n_batches)]
    pieces: Dict[tuple, List[dict]] = {}
    for f in shared:
        my_batches = [b for u, (g, b) in enumerate(units) if g == f and u % world == rank]
        if not my_batches:
            continue
        got = grid.run_layer(model, layer_of[f], by_layer[layer_of[f]], rows, max_new_tokens, temperature, batch_size,
                             layer_fraction=f, only_batches=my_batches)
        sizes = [len(part) for k, part in enumerate(grid.chunk(rows, batch_size)) if k in set(my_batches)]
        at = 0
        for b, n in zip(my_batches, sizes):
            pieces[(f, b)] = got[at:at + n]
            at += n
    if world > 1:
        parts = parallel.gather_objects_to_rank0((out, pieces))
        if rank == 0:
            out = {k: v for part, _ in parts for k, v in part.items()}
            pieces = {k: v for _, part in parts for k, v in part.items()}
    if rank == 0:
        for f in shared:
            out.update(configurations(f, [r for b in range(n_batches) for r in pieces[(f, b)]]))
        out = {(f, float(a)): out[(f, float(a))] for f in fractions for a in strengths}  # the caller's order
    return out
