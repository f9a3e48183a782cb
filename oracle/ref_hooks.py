"""ORACLE — test infrastructure, not product code.

Torch restatement of the reference's hook path: what `ModelWrapper.extract_activations`,
`ModelWrapper.generate_batch_with_multi_steering` and `vector_utils.extract_concept_vectors_batch`
compute, expressed on token ids so it can run without a tokenizer or the Hub.  It drives a stock
Hugging Face model through `register_forward_hook` + `model.generate`, exactly the mechanism the
reference uses, and therefore runs on whatever device the HF model is on: on the CPU it is the
reference's CPU path (the `cpu_baseline` / `--impl reference` leg of bench.py); on the GPU it is
the "reference hook path on the same device and dtype" that north_star's logit/greedy-token
parity is defined against.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s baseline legs may import this.

Parity status: PINNED — `oracle/make_golden.py` runs the real reference
(/root/reference/model_utils.py, vector_utils.py) and this restatement on the same seeded
models and asserts identical activations, vectors, logits and token ids before writing
`tests/golden/`; `tests/test_oracle_golden.py` re-checks this file against those fixtures.
"""

from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch


def decoder_layers(model):
    """reference model_utils.py:273-291 (`get_layer_module`): `model.model.layers`, or
    `model.model.language_model.layers` for multimodal wrappers."""
    inner = model.model
    if hasattr(inner, "layers"):
        return inner.layers
    return inner.language_model.layers


def _hidden_of(output):
    # reference model_utils.py:314-317: decoder layers return a tuple or a bare tensor.
    return output[0] if isinstance(output, tuple) else output


@torch.no_grad()
def capture(model, input_ids: torch.Tensor, attention_mask: torch.Tensor, layer_idx: int,
            token_idx: int = -1) -> torch.Tensor:
    """Residual stream after decoder layer `layer_idx` at `token_idx`, as a CPU tensor in the
    model dtype (reference model_utils.py:312-345: hook, slice, `.detach().cpu()`, full forward
    with `use_cache=False`)."""
    grabbed: List[torch.Tensor] = []

    def hook(_m, _inp, out):
        grabbed.append(_hidden_of(out)[:, token_idx, :].detach().cpu())

    handle = decoder_layers(model)[layer_idx].register_forward_hook(hook)
    try:
        dev = next(model.parameters()).device
        model(input_ids=input_ids.to(dev), attention_mask=attention_mask.to(dev), use_cache=False)
    finally:
        handle.remove()
    return torch.cat(grabbed, dim=0)


def concept_vectors(concept_acts: torch.Tensor, base_acts: Optional[torch.Tensor], method: str,
                    normalize: bool = False) -> torch.Tensor:
    """Rows of `concept_acts` turned into concept vectors, in the dtype of the activations.

    baseline:    act_c - mean(base)        reference vector_utils.py:541-548
    simple:      act_c - base[0]           reference vector_utils.py:558, 570-572
    no_baseline: act_c                     reference vector_utils.py:586-588
    posneg:      mean(concept) - mean(base) (one vector) reference vector_utils.py:103-109
    """
    if method == "baseline":
        out = concept_acts - base_acts.mean(dim=0)
    elif method == "simple":
        out = concept_acts - base_acts[0]
    elif method == "no_baseline":
        out = concept_acts.clone()
    elif method == "posneg":
        out = (concept_acts.mean(dim=0) - base_acts.mean(dim=0)).unsqueeze(0)
    else:
        raise ValueError(f"Unknown extraction method: {method}")
    if normalize:
        out = torch.stack([v / (v.norm() + 1e-8) for v in out])
    return out


@torch.no_grad()
def steered_generate(model, input_ids: torch.Tensor, attention_mask: torch.Tensor, layer_idx: int,
                     vectors: Sequence[torch.Tensor], strength: float,
                     start_positions: Optional[Sequence[int]], max_new_tokens: int,
                     temperature: float = 0.0, pad_token_id: int = 0,
                     want_logits: bool = False, tap: Optional[dict] = None) -> Dict[str, torch.Tensor]:
    """Batched generation with one steering vector per row.

    Follows reference model_utils.py:687-879: vectors are scaled by the scalar `strength` and
    cast to the model dtype once (:737); `start_positions` are shifted right by each row's
    left-pad count (:822-825); the hook clones the layer output and, when the forward sees a
    single position (a cached decode step), adds row i's vector to row i (:774-777), otherwise
    adds it to positions >= start_i of row i when start_i lies inside the sequence (:784-788);
    with no positions given every position of every row is steered (:798-802).

    Returns {"new_ids": [B, max_new_tokens]} (ids from the padded prompt width onward — the
    reference slices at the unpadded length instead, SURVEY.md A3, which only affects text) and,
    with want_logits, {"logits": [steps, B, V]} as float32 on the CPU.
    `tap`, if given, receives the hooked layer's tensors before/after the add and the next
    norm's output for the first `tap["calls"]` hook calls (used to make K2 fixtures).
    """
    dev = next(model.parameters()).device
    dtype = next(model.parameters()).dtype
    input_ids = input_ids.to(dev)
    attention_mask = attention_mask.to(dev)
    scaled = torch.stack([v * strength for v in vectors]).to(dev).to(dtype)
    if len(vectors) != input_ids.shape[0]:
        raise AssertionError("Must have one steering vector per prompt")
    starts = None
    if start_positions is not None:
        if len(start_positions) != input_ids.shape[0]:
            raise AssertionError("Must have one steering position per prompt")
        width = input_ids.shape[1]
        pads = width - attention_mask.sum(dim=1)
        starts = (torch.tensor(list(start_positions), device=dev) + pads).tolist()

    def hook(_m, _inp, out):
        hidden = _hidden_of(out)
        seq = hidden.shape[1]
        if starts is None:
            new = hidden + scaled.to(hidden.device).unsqueeze(1)
        else:
            new = hidden.clone()
            sv = scaled.to(hidden.device)
            for i in range(hidden.shape[0]):
                if seq == 1:
                    new[i, :, :] += sv[i]
                elif starts[i] < seq:
                    new[i, starts[i]:, :] += sv[i]
        if tap is not None and len(tap["before"]) < tap["calls"]:
            tap["before"].append(hidden.detach().cpu())
            tap["after"].append(new.detach().cpu())
        return (new,) + tuple(out[1:]) if isinstance(out, tuple) else new

    layers = decoder_layers(model)
    handles = [layers[layer_idx].register_forward_hook(hook)]
    if tap is not None:
        tap.setdefault("before", []); tap.setdefault("after", []); tap.setdefault("normed", [])
        nxt = layers[layer_idx + 1].input_layernorm if layer_idx + 1 < len(layers) else model.model.norm

        def norm_hook(_m, _inp, out):
            if len(tap["normed"]) < tap["calls"]:
                tap["normed"].append(out.detach().cpu())

        handles.append(nxt.register_forward_hook(norm_hook))
        tap["starts"] = starts
    kwargs = dict(max_new_tokens=max_new_tokens, pad_token_id=pad_token_id)
    if temperature > 0:
        kwargs.update(do_sample=True, temperature=temperature)
    else:
        kwargs.update(do_sample=False)
    if want_logits:
        kwargs.update(output_logits=True, return_dict_in_generate=True)
    try:
        out = model.generate(input_ids=input_ids, attention_mask=attention_mask, **kwargs)
    finally:
        for hnd in handles:
            hnd.remove()
    seqs = out.sequences if want_logits else out
    result = {"new_ids": seqs[:, input_ids.shape[1]:].cpu()}
    if want_logits:
        result["logits"] = torch.stack([l.float().cpu() for l in out.logits])
    return result


@torch.no_grad()
def hook_body_and_norm(hidden: torch.Tensor, scaled_vecs: torch.Tensor, starts: Optional[List[int]],
                       norm_module) -> torch.Tensor:
    """Path-only micro-benchmark body: the reference's clone + per-row adds
    (reference model_utils.py:767-788) followed by the next layer's HF RMSNorm."""
    new = hidden.clone()
    seq = hidden.shape[1]
    for i in range(hidden.shape[0]):
        if seq == 1:
            new[i, :, :] += scaled_vecs[i]
        elif starts is None or starts[i] < seq:
            new[i, (0 if starts is None else starts[i]):, :] += scaled_vecs[i]
    return norm_module(new)
