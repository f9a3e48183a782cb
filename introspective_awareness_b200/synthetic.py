"""Random-init models of the named architectures and seeded synthetic inputs.

There are no checkpoints offline, so every config in BASELINE.json runs on
`AutoModelForCausalLM.from_config`-style random weights.  Dimensions are the
public `config.json` values of each model; the reference only lists repo ids
(reference model_utils.py:26-44).

Weights are filled from a counter-based numpy generator (`numpy.random.Philox`), not
`torch.manual_seed`, so that the golden fixtures made in the build container and the tests on
the GPU box see bit-identical weights regardless of which torch CPU kernels the host selects.
"""

from __future__ import annotations

import zlib
from typing import Dict, Optional

import numpy as np
import torch

ARCHS: Dict[str, dict] = {
    # family, then HF config kwargs
    "qwen2.5-0.5b": dict(
        family="qwen2", vocab_size=151936, hidden_size=896, intermediate_size=4864,
        num_hidden_layers=24, num_attention_heads=14, num_key_value_heads=2,
        rms_norm_eps=1e-6, rope_theta=1000000.0, max_position_embeddings=32768,
        tie_word_embeddings=True,
    ),
    "llama-3.1-8b": dict(
        family="llama", vocab_size=128256, hidden_size=4096, intermediate_size=14336,
        num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8,
        rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
        tie_word_embeddings=False,
        rope_scaling=dict(rope_type="llama3", factor=8.0, low_freq_factor=1.0,
                          high_freq_factor=4.0, original_max_position_embeddings=8192),
    ),
    "gemma-2-9b": dict(
        family="gemma2", vocab_size=256000, hidden_size=3584, intermediate_size=14336,
        num_hidden_layers=42, num_attention_heads=16, num_key_value_heads=8, head_dim=256,
        rms_norm_eps=1e-6, rope_theta=10000.0, max_position_embeddings=8192,
        sliding_window=4096, query_pre_attn_scalar=256,
        attn_logit_softcapping=50.0, final_logit_softcapping=30.0,
    ),
    "qwen2.5-32b": dict(
        family="qwen2", vocab_size=152064, hidden_size=5120, intermediate_size=27648,
        num_hidden_layers=64, num_attention_heads=40, num_key_value_heads=8,
        rms_norm_eps=1e-6, rope_theta=1000000.0, max_position_embeddings=32768,
        tie_word_embeddings=False,
    ),
    # Small stand-ins with the same code paths, for tests and smoke().
    "tiny-qwen2": dict(
        family="qwen2", vocab_size=512, hidden_size=128, intermediate_size=256,
        num_hidden_layers=4, num_attention_heads=4, num_key_value_heads=2,
        rms_norm_eps=1e-6, rope_theta=10000.0, max_position_embeddings=512,
        tie_word_embeddings=True,
    ),
    "tiny-llama": dict(
        family="llama", vocab_size=512, hidden_size=128, intermediate_size=256,
        num_hidden_layers=4, num_attention_heads=4, num_key_value_heads=2,
        rms_norm_eps=1e-5, rope_theta=10000.0, max_position_embeddings=512,
        tie_word_embeddings=False,
    ),
    "tiny-gemma2": dict(
        family="gemma2", vocab_size=512, hidden_size=128, intermediate_size=256,
        num_hidden_layers=4, num_attention_heads=4, num_key_value_heads=2, head_dim=32,
        rms_norm_eps=1e-6, rope_theta=10000.0, max_position_embeddings=512,
        sliding_window=64, query_pre_attn_scalar=32,
        attn_logit_softcapping=50.0, final_logit_softcapping=30.0,
    ),
    # Real hidden size and vocabulary of Llama-3.1-8B but only 4 layers: K2 sees the true
    # [rows, 4096] shapes while the model builds in seconds.
    "llama-8b-4layer": dict(
        family="llama", vocab_size=128256, hidden_size=4096, intermediate_size=14336,
        num_hidden_layers=4, num_attention_heads=32, num_key_value_heads=8,
        rms_norm_eps=1e-5, rope_theta=500000.0, max_position_embeddings=131072,
        tie_word_embeddings=False,
        rope_scaling=dict(rope_type="llama3", factor=8.0, low_freq_factor=1.0,
                          high_freq_factor=4.0, original_max_position_embeddings=8192),
    ),
    # Real Gemma-2-9B and Qwen2.5-32B widths / head layouts / vocabularies with few layers, for the
    # same-device parity tests of BASELINE configs[3] and [4] (Gemma: 4 norms per layer, 1 + w scaling,
    # soft-capping, sliding window; Qwen: H = 5120, q/k/v biases).
    "gemma-2-9b-4layer": dict(
        family="gemma2", vocab_size=256000, hidden_size=3584, intermediate_size=14336,
        num_hidden_layers=4, num_attention_heads=16, num_key_value_heads=8, head_dim=256,
        rms_norm_eps=1e-6, rope_theta=10000.0, max_position_embeddings=8192,
        sliding_window=4096, query_pre_attn_scalar=256,
        attn_logit_softcapping=50.0, final_logit_softcapping=30.0,
    ),
    "qwen2.5-32b-2layer": dict(
        family="qwen2", vocab_size=152064, hidden_size=5120, intermediate_size=27648,
        num_hidden_layers=2, num_attention_heads=40, num_key_value_heads=8,
        rms_norm_eps=1e-6, rope_theta=1000000.0, max_position_embeddings=32768,
        tie_word_embeddings=False,
    ),
}


def make_config(arch: str, **overrides):
    """HF config object for a named architecture (no weights)."""
    from transformers import Gemma2Config, LlamaConfig, Qwen2Config

    spec = dict(ARCHS[arch])
    spec.update(overrides)
    family = spec.pop("family")
    cls = {"qwen2": Qwen2Config, "llama": LlamaConfig, "gemma2": Gemma2Config}[family]
    rope_scaling = spec.pop("rope_scaling", None)
    rope_theta = spec.pop("rope_theta", None)
    cfg = cls(**spec)
    # transformers 5 keeps rope settings in `rope_parameters`; 4.x uses rope_theta/rope_scaling.
    if hasattr(cfg, "rope_parameters"):
        params = dict(rope_scaling or {"rope_type": "default"})
        if rope_theta is not None:
            params["rope_theta"] = rope_theta
        cfg.rope_parameters = params
    else:  # pragma: no cover - older transformers
        if rope_theta is not None:
            cfg.rope_theta = rope_theta
        if rope_scaling is not None:
            cfg.rope_scaling = rope_scaling
    # No early stop on random weights unless a test asks for it.
    cfg.eos_token_id = None
    cfg.pad_token_id = 0
    cfg.bos_token_id = None
    return cfg


def seeded_fill_(model: torch.nn.Module, seed: int = 0, std: float = 0.02) -> None:
    """Overwrite every parameter with platform-independent seeded values.

    Matrices ~ N(0, std); RMSNorm weights ~ N(1, 0.02) (Gemma: N(0, 0.02), its norm
    multiplies by 1 + w); biases ~ N(0, 0.02).  One Philox stream per parameter name so
    adding a parameter never shifts the others.
    """
    is_gemma = "gemma" in type(model).__name__.lower()
    with torch.no_grad():
        seen = set()
        for name, p in model.named_parameters():
            if p.data_ptr() in seen:  # tied weights
                continue
            seen.add(p.data_ptr())
            key = (zlib.crc32(name.encode()) << 20) ^ int(seed)
            rng = np.random.Generator(np.random.Philox(key=key))
            vals = rng.standard_normal(p.numel(), dtype=np.float32)
            t = torch.from_numpy(vals).reshape(p.shape)
            if "norm" in name and p.dim() == 1:
                t = t * 0.02 + (0.0 if is_gemma else 1.0)
            else:
                t = t * std
            p.copy_(t.to(p.dtype))


def build_model(arch: str, dtype: torch.dtype = torch.float32, device: str = "cpu",
                seed: Optional[int] = 0, attn_implementation: str = "sdpa", **overrides):
    """Random-init HF causal LM of `arch` in eval mode.

    `seed=None` leaves HF's own initialisation (fast on GPU; used by bench.py where only
    shapes matter); an int seed applies `seeded_fill_` on the CPU and then moves the model.
    """
    from transformers import AutoModelForCausalLM

    cfg = make_config(arch, **overrides)
    cfg._attn_implementation = attn_implementation
    if seed is None:
        prev = torch.get_default_dtype()
        torch.set_default_dtype(dtype)
        try:
            with torch.device(device):
                model = AutoModelForCausalLM.from_config(cfg, attn_implementation=attn_implementation)
        finally:
            torch.set_default_dtype(prev)
        model = model.to(device=device, dtype=dtype)
    else:
        model = AutoModelForCausalLM.from_config(cfg, attn_implementation=attn_implementation)
        model = model.to(torch.float32)
        seeded_fill_(model, seed)
        model = model.to(device=device, dtype=dtype)
    if getattr(model, "generation_config", None) is not None:
        model.generation_config.eos_token_id = None
        model.generation_config.pad_token_id = 0
    return model.eval()


def fast_random_fill_(model: torch.nn.Module, seed: int = 0, std: float = 0.02) -> None:
    """On-device N(0, std) fill for big bench models (values need not be reproducible
    across hosts, only well-scaled)."""
    gen = torch.Generator(device=next(model.parameters()).device)
    gen.manual_seed(seed)
    is_gemma = "gemma" in type(model).__name__.lower()
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "norm" in name and p.dim() == 1:
                p.normal_(0.0 if is_gemma else 1.0, 0.02, generator=gen)
            else:
                p.normal_(0.0, std, generator=gen)


def synthetic_prompt_ids(n_rows: int, t_max: int, t_min: int, vocab_size: int, seed: int = 1,
                         pad_id: int = 0):
    """Left-padded random token ids with per-row real length U{t_min..t_max}."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    lens = rng.integers(t_min, t_max + 1, size=n_rows)
    lens[0] = t_max  # at least one full-width row so the padded width is t_max
    ids = np.full((n_rows, t_max), pad_id, dtype=np.int64)
    mask = np.zeros((n_rows, t_max), dtype=np.int64)
    for i, n in enumerate(lens):
        ids[i, t_max - n:] = rng.integers(2, vocab_size, size=n)
        mask[i, t_max - n:] = 1
    return torch.from_numpy(ids), torch.from_numpy(mask), torch.from_numpy(lens.astype(np.int64))
