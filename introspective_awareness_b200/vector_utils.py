"""Concept-vector extraction with the reference's call surface (reference vector_utils.py), the
arithmetic done by the fused K1 kernel instead of CPU tensor ops.

Every public name of the reference module exists here with the same positional order,
defaults and return types; vectors come back as detached CPU tensors in the model dtype, as the
reference's do.  `model` is our `ModelWrapper`; any other object that offers
`extract_activations(prompts, layer_idx, token_idx)` is also accepted — its activations are then
shipped to the GPU and finished by the K1b kernel (still no CPU arithmetic).

New here (north_star: capture all layers in one forward):
`extract_concept_vectors_multi_layer`.
"""

from __future__ import annotations

import json
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .model_utils import ModelWrapper, get_layer_at_fraction  # noqa: F401  (re-exported like the reference)

DEFAULT_TEMPLATE = "Tell me about {word}"

# The paper's 100 baseline words, in the reference's order (vector_utils.py:384-405;
# "Butterflies" appears twice there and is kept so that the mean is the same).
DEFAULT_BASELINE_WORDS = (
    "Desks Jackets Gondolas Laughter Intelligence Bicycles Chairs Orchestras Sand Pottery "
    "Arrowheads Jewelry Daffodils Plateaus Estuaries Quilts Moments Bamboo Ravines Archives "
    "Hieroglyphs Stars Clay Fossils Wildlife Flour Traffic Bubbles Honey Geodes "
    "Magnets Ribbons Zigzags Puzzles Tornadoes Anthills Galaxies Poverty Diamonds Universes "
    "Vinegar Nebulae Knowledge Marble Fog Rivers Scrolls Silhouettes Marbles Cakes "
    "Valleys Whispers Pendulums Towers Tables Glaciers Whirlpools Jungles Wool Anger "
    "Ramparts Flowers Research Hammers Clouds Justice Dogs Butterflies Needles Fortresses "
    "Bonfires Skyscrapers Caravans Patience Bacon Velocities Smoke Electricity Sunsets Anchors "
    "Parchments Courage Statues Oxygen Time Butterflies Fabric Pasta Snowflakes Mountains "
    "Echoes Pianos Sanctuaries Abysses Air Dewdrops Gardens Literature Rice Enigmas"
).split()

CONCEPT_PAIRS = {
    "all_caps": ("HI! HOW ARE YOU? I'M DOING GREAT TODAY!", "Hi! How are you? I'm doing great today!"),
    "recursion_code": (
        "def factorial(n):\n    if n <= 1:\n        return 1\n    return n * factorial(n - 1)",
        "def factorial(n):\n    result = 1\n    for i in range(2, n + 1):\n        result *= i\n    return result",
    ),
    "if_statement_code": (
        "def check_positive(x):\n    if x > 0:\n        return True\n    return False",
        "def check_positive(x):\n    result = x > 0\n    return result",
    ),
    "loop_code": ("for i in range(10):\n    print(i)", "print(list(range(10)))"),
}


# -- prompt helpers ---------------------------------------------------------------------------

def _chat_prompt(model, user_message: str) -> str:
    """One user turn + generation prompt; plain-text fallback when the tokenizer has no chat
    template (reference vector_utils.py:506-517)."""
    tok = model.tokenizer
    if hasattr(tok, "apply_chat_template"):
        return tok.apply_chat_template([{"role": "user", "content": user_message}], tokenize=False,
                                       add_generation_prompt=True)
    return f"User: {user_message}\n\nAssistant:"


def create_concept_prompts(concept_word: str, baseline_words: Optional[List[str]] = None,
                           template: str = DEFAULT_TEMPLATE) -> Tuple[List[str], List[str]]:
    """reference vector_utils.py:20-43"""
    return [template.format(word=concept_word)], [template.format(word=w) for w in (baseline_words or [])]


def create_contrastive_pair(positive_text: str, negative_text: str) -> Tuple[str, str]:
    return positive_text, negative_text


def get_baseline_words(n: int = 100) -> List[str]:
    return DEFAULT_BASELINE_WORDS[:n]


def get_concept_pair(concept_name: str) -> Tuple[str, str]:
    if concept_name not in CONCEPT_PAIRS:
        raise ValueError(f"Unknown concept pair: {concept_name}. Available: {list(CONCEPT_PAIRS.keys())}")
    return CONCEPT_PAIRS[concept_name]


# -- the kernel-backed core -------------------------------------------------------------------

_METHODS = ("baseline", "simple", "no_baseline", "posneg")


def _vectors(model, concept_prompts: List[str], base_prompts: Optional[List[str]], layer_indices: Sequence[int],
             method: str, normalize: bool, token_idx: int, sharded: bool = False) -> torch.Tensor:
    """[n_layers, n_out, H] CPU tensor in the model dtype."""
    if method not in _METHODS:
        raise ValueError(f"Unknown extraction method: {method}")
    if hasattr(model, "extract_concept_vectors"):
        if sharded:
            return model.extract_concept_vectors(concept_prompts, base_prompts, layer_indices, method, normalize, token_idx,
                                                 sharded=True)
        return model.extract_concept_vectors(concept_prompts, base_prompts, layer_indices, method, normalize, token_idx)
    # Foreign model object: take its activations, finish on the GPU with K1b.
    from . import kernels

    if not torch.cuda.is_available():
        raise RuntimeError("concept-vector arithmetic runs in the CUDA kernels; there is no CPU path")
    outs = []
    for layer in layer_indices:
        cact = model.extract_activations(prompts=concept_prompts, layer_idx=layer, token_idx=token_idx).cuda()
        if method == "no_baseline":
            _, vr, _ = kernels.meandiff_finalize(cact[None].contiguous(), None, 0, "no_baseline", normalize, want_f32=False)
        else:
            bact = model.extract_activations(prompts=base_prompts, layer_idx=layer, token_idx=token_idx).cuda()
            if method == "posneg":
                _, vr, _ = kernels.gather_meandiff([cact[:, None].contiguous()], [bact[:, None].contiguous()], "posneg",
                                                   normalize, want_f32=False)
            else:
                _, bsum = kernels.gather_segsum([bact[:, None].contiguous()], want_rows=False)
                _, vr, _ = kernels.meandiff_finalize(cact[None].contiguous(), bsum, bact.shape[0], method, normalize,
                                                     want_f32=False)
        outs.append(vr[0].cpu())
    return torch.stack(outs)


# -- reference call surface -------------------------------------------------------------------

def extract_concept_vector(model, positive_prompts: List[str], negative_prompts: List[str], layer_idx: int,
                           token_idx: int = -1, normalize: bool = False) -> torch.Tensor:
    """mean(positive) - mean(negative) at one layer (reference vector_utils.py:63-111)."""
    return _vectors(model, positive_prompts, negative_prompts, [layer_idx], "posneg", normalize, token_idx)[0, 0]


def extract_concept_vector_with_baseline(model, concept_word: str, baseline_words: List[str], layer_idx: int,
                                         template: str = DEFAULT_TEMPLATE, token_idx: int = -1,
                                         normalize: bool = False) -> torch.Tensor:
    """activation(concept) - mean(activation(baseline words)) (reference vector_utils.py:114-183)."""
    c = [_chat_prompt(model, template.format(word=concept_word))]
    b = [_chat_prompt(model, template.format(word=w)) for w in baseline_words]
    return _vectors(model, c, b, [layer_idx], "baseline", normalize, token_idx)[0, 0]


def extract_concept_vector_simple(model, concept_word: str, layer_idx: int, control_prompt: str = "The",
                                  template: str = "{word}", token_idx: int = -1, normalize: bool = False) -> torch.Tensor:
    """activation(concept) - activation(single control prompt) (reference vector_utils.py:186-251)."""
    c = [_chat_prompt(model, template.format(word=concept_word))]
    b = [_chat_prompt(model, control_prompt)]
    return _vectors(model, c, b, [layer_idx], "simple", normalize, token_idx)[0, 0]


def extract_concept_vector_no_baseline(model, concept_word: str, layer_idx: int, template: str = "{word}",
                                       token_idx: int = -1, normalize: bool = False) -> torch.Tensor:
    """raw activation(concept) (reference vector_utils.py:254-307)."""
    c = [_chat_prompt(model, template.format(word=concept_word))]
    return _vectors(model, c, None, [layer_idx], "no_baseline", normalize, token_idx)[0, 0]


def extract_concept_vectors_multi_layer(model, concept_words: List[str], baseline_words: List[str],
                                        layer_indices: Sequence[int], extraction_method: str = "baseline",
                                        template: str = DEFAULT_TEMPLATE, token_idx: int = -1,
                                        normalize: bool = False, sharded: bool = False) -> Dict[int, Dict[str, torch.Tensor]]:
    """All requested layers from ONE forward per prompt batch (the reference re-runs the forward for
    every layer fraction, detect_injected_thoughts.py:1551-1561).  sharded=True: a collective call of a
    `torch.distributed` job, the prompts split over its ranks (ModelWrapper.extract_vectors_sharded_ids)."""
    if extraction_method not in ("baseline", "simple", "no_baseline"):
        raise ValueError(f"Unknown extraction method: {extraction_method}")
    if not concept_words:  # the reference's loops simply produce an empty dict
        return {int(layer): {} for layer in layer_indices}
    concept_prompts = [_chat_prompt(model, template.format(word=w)) for w in concept_words]
    if extraction_method == "baseline":
        base_prompts = [_chat_prompt(model, template.format(word=w)) for w in baseline_words]
    elif extraction_method == "simple":
        base_prompts = [_chat_prompt(model, template.format(word="The"))]  # reference :552-553
    else:
        base_prompts = None
    vecs = _vectors(model, concept_prompts, base_prompts, list(layer_indices), extraction_method, normalize, token_idx,
                    sharded=sharded)
    return {int(layer): {w: vecs[li, ci] for ci, w in enumerate(concept_words)}
            for li, layer in enumerate(layer_indices)}


def extract_concept_vectors_batch(model, concept_words: List[str], baseline_words: List[str], layer_idx: int,
                                  extraction_method: str = "baseline", template: str = DEFAULT_TEMPLATE,
                                  token_idx: int = -1, normalize: bool = False) -> Dict[str, torch.Tensor]:
    """reference vector_utils.py:480-594: Dict[word -> Tensor[H]]."""
    return extract_concept_vectors_multi_layer(model, concept_words, baseline_words, [layer_idx], extraction_method,
                                               template, token_idx, normalize)[int(layer_idx)]


def cosine_similarity(vec1: torch.Tensor, vec2: torch.Tensor) -> float:
    """<a,b> / (|a||b| + 1e-8) in the vectors' own dtype (reference vector_utils.py:310-328)."""
    a, b = vec1.flatten(), vec2.flatten()
    return ((a * b).sum() / (a.norm() * b.norm() + 1e-8)).item()


def save_concept_vector(vector: torch.Tensor, save_path: Path, metadata: Optional[Dict] = None):
    """Bare `torch.save` of the tensor plus optional JSON side-car (reference vector_utils.py:331-356)."""
    save_path = Path(save_path)
    save_path.parent.mkdir(parents=True, exist_ok=True)
    torch.save(vector, save_path)
    if metadata is not None:
        with open(save_path.with_suffix(".json"), "w") as f:
            json.dump(metadata, f, indent=2)
    print(f"Saved concept vector to {save_path}")


def load_concept_vector(load_path: Path) -> Tuple[torch.Tensor, Optional[Dict]]:
    load_path = Path(load_path)
    vector = torch.load(load_path)
    meta_path = load_path.with_suffix(".json")
    metadata = json.load(open(meta_path)) if meta_path.exists() else None
    return vector, metadata


def analyze_vector_underspecification(model, target_concept: str, related_concepts: List[str], layer_idx: int,
                                      baseline_words: Optional[List[str]] = None) -> Dict[str, float]:
    """Cosine similarity of the target vector with each related concept's vector
    (reference vector_utils.py:597-643); all vectors come from one batched extraction."""
    baseline_words = baseline_words if baseline_words is not None else get_baseline_words()
    words = [target_concept] + [c for c in related_concepts]
    vecs = extract_concept_vectors_batch(model, list(dict.fromkeys(words)), baseline_words, layer_idx)
    return {c: cosine_similarity(vecs[target_concept], vecs[c]) for c in related_concepts}
