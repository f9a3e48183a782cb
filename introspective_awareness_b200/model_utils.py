"""`ModelWrapper` with the reference's method surface (reference model_utils.py:56-900), driving
the B200 runtime (engine.py) instead of forward hooks + `model.generate`.

Same names, argument order, defaults and return types as the reference so that
`vector_utils`, `steering_utils` and the reference driver can call it unchanged:

    extract_activations(prompts, layer_idx, token_idx=-1) -> Tensor[B, H]   (CPU, model dtype)
    generate_with_steering(prompt, layer_idx, steering_vector, strength=1.0, max_new_tokens=512,
                           temperature=0.0, steering_start_pos=None, **kw) -> str
    generate(prompt, ...) -> str, generate_batch(prompts, ...) -> List[str]
    generate_batch_with_steering(prompts, layer_idx, steering_vector, ...) -> List[str]
    generate_batch_with_multi_steering(prompts, layer_idx, steering_vectors, strength=1.0, ...,
                                       steering_start_positions=None, debug=False, **kw) -> List[str]
    get_layer_module(layer_idx), cleanup(), n_layers, tokenizer, device, dtype, model_name

Where the reference's three steering entry points disagree this class
implements ONE semantics, that of `generate_batch_with_multi_steering`, the only one the sweep
uses: prompt positions >= t0 (shifted by the row's left padding) and every generated token.
  * A1  `generate_with_steering` steers generated tokens too (the reference's hook does not on
        cached steps, and not at all when layers return bare tensors);
  * A2  `generate_batch_with_steering` works (the reference raises UnboundLocalError at :643);
  * A3  returned text starts at the first NEW token; set `compat_slice=True` to reproduce the
        reference's strings, which begin `pad_i` prompt tokens early (:870-871).
Extensions (not in the reference): `strength` may be a per-row sequence (north_star: a different
(concept, alpha, t0) per batch row), `steering_vectors` may repeat the same tensor object (rows
share one table entry), and `extract_activations_multi` / `extract_concept_vectors` capture any
number of layers in one forward.
"""

from __future__ import annotations

import gc
from typing import Dict, List, Optional, Sequence, Union

import torch

from .engine import DecoderEngine, SteerSpec

# Same registry as the reference (model_utils.py:19-47): friendly name -> Hugging Face repo id.
MODEL_NAME_MAP = {
    "llama_8b": "meta-llama/Llama-3.1-8B-Instruct",
    "llama_70b": "meta-llama/Llama-3.3-70B-Instruct",
    "llama_405b": "meta-llama/Llama-3.1-405B-Instruct",
    "qwen_7b": "Qwen/Qwen2.5-7B-Instruct",
    "qwen_14b": "Qwen/Qwen2.5-14B-Instruct",
    "qwen_32b": "Qwen/Qwen2.5-32B-Instruct",
    "qwen_72b": "Qwen/Qwen2.5-72B-Instruct",
    "gemma2_2b": "google/gemma-2-2b-it",
    "gemma2_9b": "google/gemma-2-9b-it",
    "gemma2_27b": "google/gemma-2-27b-it",
}


class ModelWrapper:
    def __init__(self, model_name: str, device: str = "cuda", dtype: torch.dtype = torch.bfloat16,
                 quantization_config=None, *, model=None, tokenizer=None, compat_slice: bool = False,
                 attention: str = "auto"):
        self.model_name = model_name
        self.device = device
        self.dtype = dtype
        self.hf_path = MODEL_NAME_MAP.get(model_name, model_name)
        self.compat_slice = compat_slice
        if model is None:
            # Stock Hugging Face loading, as reference model_utils.py:91-126 (needs the Hub).
            from transformers import AutoModelForCausalLM, AutoTokenizer

            if quantization_config is not None:
                raise NotImplementedError("bitsandbytes quantisation is outside the B200 hot path")
            tokenizer = tokenizer or AutoTokenizer.from_pretrained(self.hf_path)
            model = AutoModelForCausalLM.from_pretrained(self.hf_path, dtype=dtype).to(device)
        self.model = model.eval()
        self.tokenizer = tokenizer
        if tokenizer is not None:
            tokenizer.padding_side = "left"  # reference model_utils.py:97
            if getattr(tokenizer, "pad_token", None) is None:
                tokenizer.pad_token = tokenizer.eos_token
        self.engine = DecoderEngine(self.model, attention=attention)
        self.n_layers = self.engine.n_layers
        self.hooks: List = []  # kept for interface parity; this runtime registers no hooks

    # -- construction helpers -----------------------------------------------------------------
    @classmethod
    def from_arch(cls, arch: str, device: str = "cuda", dtype: torch.dtype = torch.bfloat16, seed: Optional[int] = 0,
                  tokenizer=None, attention: str = "auto", **overrides) -> "ModelWrapper":
        """Random-init model of a named architecture + stub tokenizer (offline)."""
        from .stub_tokenizer import StubTokenizer
        from .synthetic import build_model

        model = build_model(arch, dtype=dtype, device=device, seed=seed, **overrides)
        tok = tokenizer or StubTokenizer(vocab_size=model.config.vocab_size)
        return cls(arch, device=device, dtype=dtype, model=model, tokenizer=tok, attention=attention)

    def _get_input_device(self):
        return next(self.model.parameters()).device

    def _get_n_layers(self) -> int:
        return self.engine.n_layers

    def get_layer_module(self, layer_idx: int):
        return self.engine.layers[layer_idx]

    # -- tokenisation -------------------------------------------------------------------------
    def _encode(self, prompts: Sequence[str]):
        enc = self.tokenizer(list(prompts), return_tensors="pt", padding=True, truncation=True)
        return enc["input_ids"], enc["attention_mask"]

    def _eos(self):
        """Every id of the checkpoint's EOS set (generation_config.eos_token_id may be a list: Llama-3.x-Instruct
        [128001, 128008, 128009], Gemma-2-it [1, 107]); HF `generate` ends a row on any of them."""
        gc_ = getattr(self.model, "generation_config", None)
        eos = getattr(gc_, "eos_token_id", None) if gc_ is not None else None
        if isinstance(eos, (list, tuple)):
            eos = tuple(int(e) for e in eos) or None
        return eos

    def _generation_settings(self, temperature: float, generation_kwargs: dict) -> dict:
        """What reference model_utils.py:828-847 hands to HF `generate`, resolved the way `generate` resolves it:
        the caller's temperature switches sampling on and carries **generation_kwargs (:833-837); everything it does
        not set comes from the checkpoint's generation_config — top_k / top_p / repetition_penalty, and with
        temperature == 0 also do_sample and the temperature themselves (the reference passes neither then, so a
        checkpoint that ships do_sample=True still samples)."""
        gc_ = getattr(self.model, "generation_config", None)
        cfg = lambda k, d: d if gc_ is None or getattr(gc_, k, None) is None else getattr(gc_, k)
        if temperature > 0:
            over = dict(generation_kwargs)
            do_sample, temp = bool(over.get("do_sample", True)), float(over.get("temperature", temperature))
        else:
            over = {}
            do_sample, temp = bool(cfg("do_sample", False)), float(cfg("temperature", 1.0))
        pick = lambda k, d: over[k] if k in over else cfg(k, d)
        return dict(do_sample=do_sample, temperature=temp if do_sample else 0.0, top_k=pick("top_k", 50),
                    top_p=pick("top_p", 1.0), repetition_penalty=float(pick("repetition_penalty", 1.0)))

    def _decode_rows(self, new_ids: torch.Tensor, input_ids: torch.Tensor, attention_mask: torch.Tensor) -> List[str]:
        new_ids = new_ids.cpu()
        out = []
        width = input_ids.shape[1]
        lens = attention_mask.sum(dim=1).tolist()
        for i in range(new_ids.shape[0]):
            row = new_ids[i]
            if self.compat_slice and lens[i] < width:
                # reference model_utils.py:870-871 slices at the unpadded length
                row = torch.cat([input_ids[i, lens[i]:].cpu(), row])
            out.append(self.tokenizer.decode(row, skip_special_tokens=True).strip())
        return out

    # -- extraction (reference model_utils.py:293-345) ----------------------------------------
    def extract_activations(self, prompts: List[str], layer_idx: int, token_idx: int = -1) -> torch.Tensor:
        return self.extract_activations_multi(prompts, [layer_idx], token_idx)[0]

    def extract_activations_multi(self, prompts: List[str], layer_indices: Sequence[int], token_idx: int = -1
                                  ) -> torch.Tensor:
        """[n_layers, B, H] CPU tensor in the model dtype: one forward for all layers."""
        if self.engine.device.type != "cuda":
            raise RuntimeError("activation capture runs inside the CUDA kernels (K2's tap); there is no CPU path")
        ids, mask = self._encode(prompts)
        return self.engine.capture_rows(ids, mask, layer_indices, token_idx).cpu()

    def extract_concept_vectors(self, concept_prompts: List[str], baseline_prompts: Optional[List[str]],
                                layer_indices: Sequence[int], method: str = "baseline", normalize: bool = False,
                                token_idx: int = -1, fp32: bool = False, sharded: bool = False) -> torch.Tensor:
        """Fused K1 path: [n_layers, n_out, H] CPU tensor — the reference-rounded vectors in the model
        dtype, or with fp32=True the un-rounded fp32-accumulate vectors.
        sharded=True (a collective call: every rank of the `torch.distributed` job must make it with the same
        arguments) splits the prompts over the ranks, see `extract_concept_vectors_sharded`."""
        cids, cmask = self._encode(concept_prompts)
        bids = bmask = None
        if method not in ("no_baseline", "raw"):
            bids, bmask = self._encode(baseline_prompts)
        if sharded:
            v32, vr, _ = self.extract_vectors_sharded_ids(cids, cmask, bids, bmask, layer_indices, method, normalize,
                                                          token_idx, want_f32=fp32)
        else:
            v32, vr, _ = self.engine.extract_vectors(cids, cmask, bids, bmask, layer_indices, mode=method,
                                                     normalize=normalize, token_idx=token_idx, want_f32=fp32, want_norm=False)
        return (v32 if fp32 else vr).cpu()

    def extract_vectors_sharded_ids(self, concept_ids, concept_mask, base_ids, base_mask, layer_indices: Sequence[int],
                                    method: str = "baseline", normalize: bool = False, token_idx: int = -1,
                                    want_f32: bool = False):
        """Prompt-sharded K1: every rank forwards its contiguous slice of the concept prompts and of
        the baseline prompts (captured rows land compactly through K2's tap), K1a sums its baseline rows in fp32,
        the [L, H] sums are all-reduced and the concept rows all-gathered (NCCL; gloo on CPU test rigs), and K1b
        finishes on every rank — rounding to the model dtype happens after the reduce.  Token-level entry point;
        returns (v_f32 | None, v_rounded, norm) on the device, identical on every rank.  The slices keep the padded
        width of the whole batch, so every row sees the positions it would see un-sharded."""
        from . import kernels, parallel

        if method == "posneg":
            raise ValueError("pos/neg mean difference is not sharded: call the un-sharded path")
        layer_indices = [int(i) for i in layer_indices]
        eng = self.engine
        n_concept = concept_ids.shape[0]
        n_base = 0 if base_ids is None else base_ids.shape[0]

        def capture_rows_and_sum(segment, lo, hi):
            ids, mask = (concept_ids, concept_mask) if segment == "concept" else (base_ids, base_mask)
            if hi > lo:
                rows = eng.capture_rows(ids[lo:hi], mask[lo:hi], layer_indices, token_idx, slot=0 if segment == "concept" else 1)
            else:  # more ranks than prompts: this rank contributes nothing to this segment
                rows = torch.empty((len(layer_indices), 0, eng.hidden), dtype=eng.dtype, device=eng.device)
            if segment == "concept":
                return rows, None
            eng.counters["k1_launches"] += 1
            return None, kernels.segsum_compact(rows)

        def finalize(rows, base_sum, n_base_total, mode, norm):
            eng.counters["k1_launches"] += 2 if norm else 1
            return kernels.meandiff_finalize(rows.contiguous(), base_sum, n_base_total, mode, norm, want_f32=want_f32)

        return parallel.extract_vectors_sharded(capture_rows_and_sum, finalize, n_concept, n_base, method, normalize)

    # -- generation ---------------------------------------------------------------------------
    def steered_generate_ids(self, input_ids: torch.Tensor, attention_mask: torch.Tensor, layer_idx: Optional[int],
                             vtable: Optional[torch.Tensor], row_vec, row_alpha, start_positions,
                             max_new_tokens: int, temperature: float = 0.0, keep_logits: bool = False,
                             generator: Optional[torch.Generator] = None, **generation_kwargs) -> torch.Tensor:
        """Token-level entry point.  `start_positions` are the reference's un-shifted t0 values
        (or None for all positions); the left-pad shift of reference model_utils.py:822-825 is applied
        here.  Returns new ids [B, max_new_tokens] on the device."""
        b, t = input_ids.shape
        dev = self.engine.device
        steer = None
        if layer_idx is not None and vtable is not None:
            if dev.type != "cuda":
                raise RuntimeError("steered generation needs the CUDA kernels; there is no CPU path")
            rv = torch.as_tensor(row_vec, dtype=torch.int64)
            if rv.numel() != b or (rv.numel() and int(rv.max()) >= int(vtable.shape[0])):
                raise IndexError(f"row_vec must hold {b} indices below the table's {int(vtable.shape[0])} rows "
                                 "(negative = control row)")
            t0 = None
            if start_positions is not None:
                pads = t - attention_mask.sum(dim=1)
                t0 = (torch.as_tensor(start_positions, dtype=torch.int64) + pads.cpu()).to(torch.int32)
            steer = SteerSpec(
                int(layer_idx), vtable.to(dev, non_blocking=True),
                torch.as_tensor(row_vec, dtype=torch.int32), torch.as_tensor(row_alpha, dtype=torch.float32),
                t0)
        padded = not bool(attention_mask.all())
        n_vec_cap = 64 if vtable is None else max(64, int(vtable.shape[0]))
        sess = self.engine.session(b, t, max_new_tokens, padded, n_vec_cap=n_vec_cap,
                                   vec_dtype=None if vtable is None else vtable.dtype)
        self.last_session = sess
        gs = self._generation_settings(temperature, generation_kwargs)
        return sess.generate(input_ids, attention_mask, steer, temperature=gs["temperature"], top_k=gs["top_k"],
                             top_p=gs["top_p"], repetition_penalty=gs["repetition_penalty"], do_sample=gs["do_sample"],
                             eos_token_id=self._eos(),
                             pad_token_id=self.tokenizer.pad_token_id if self.tokenizer is not None else 0,
                             generator=generator, keep_logits=keep_logits)

    def _vector_table(self, vectors: Sequence[torch.Tensor]):
        """De-duplicate by tensor identity: rows that share a concept share a table entry."""
        index: Dict[int, int] = {}
        uniq: List[torch.Tensor] = []
        row_vec = []
        for v in vectors:
            k = id(v)
            if k not in index:
                index[k] = len(uniq)
                uniq.append(v.detach().reshape(-1))
            row_vec.append(index[k])
        table = torch.stack(uniq)
        if table.dtype not in (torch.float32, self.dtype):
            table = table.to(torch.float32)
        return table, row_vec

    def generate_batch_with_multi_steering(self, prompts: List[str], layer_idx: int,
                                           steering_vectors: List[torch.Tensor],
                                           strength: Union[float, Sequence[float]] = 1.0, max_new_tokens: int = 512,
                                           temperature: float = 0.0,
                                           steering_start_positions: Optional[List[int]] = None, debug: bool = False,
                                           **generation_kwargs) -> List[str]:
        assert len(prompts) == len(steering_vectors), "Must have one steering vector per prompt"
        if not prompts:
            return []
        if steering_start_positions is not None:
            assert len(steering_start_positions) == len(prompts), "Must have one steering position per prompt"
        ids, mask = self._encode(prompts)
        table, row_vec = self._vector_table(steering_vectors)
        alphas = [float(strength)] * len(prompts) if not isinstance(strength, (list, tuple)) else [float(a) for a in strength]
        assert len(alphas) == len(prompts), "Must have one strength per prompt"
        new_ids = self.steered_generate_ids(ids, mask, layer_idx, table, row_vec, alphas, steering_start_positions,
                                            max_new_tokens, temperature, **generation_kwargs)
        if debug:
            n = self.engine.counters["k2_launches"]
            print(f"[DEBUG] fused inject+RMSNorm launches so far: {n} (1 prompt pass + {max_new_tokens - 1} decode steps per call)")
        return self._decode_rows(new_ids, ids, mask)

    def generate_batch_with_steering(self, prompts: List[str], layer_idx: int, steering_vector: torch.Tensor,
                                     strength: float = 1.0, max_new_tokens: int = 512, temperature: float = 0.0,
                                     steering_start_pos: Optional[int] = None, **generation_kwargs) -> List[str]:
        starts = None if steering_start_pos is None else [steering_start_pos] * len(prompts)
        return self.generate_batch_with_multi_steering(prompts, layer_idx, [steering_vector] * len(prompts), strength,
                                                       max_new_tokens, temperature, starts, **generation_kwargs)

    def generate_with_steering(self, prompt: str, layer_idx: int, steering_vector: torch.Tensor, strength: float = 1.0,
                               max_new_tokens: int = 512, temperature: float = 0.0,
                               steering_start_pos: Optional[int] = None, **generation_kwargs) -> str:
        starts = None if steering_start_pos is None else [steering_start_pos]
        return self.generate_batch_with_multi_steering([prompt], layer_idx, [steering_vector], strength, max_new_tokens,
                                                       temperature, starts, **generation_kwargs)[0]

    def generate_batch(self, prompts: List[str], max_new_tokens: int = 512, temperature: float = 0.0,
                       **generation_kwargs) -> List[str]:
        if not prompts:
            return []
        ids, mask = self._encode(prompts)
        new_ids = self.steered_generate_ids(ids, mask, None, None, None, None, None, max_new_tokens, temperature,
                                            **generation_kwargs)
        return self._decode_rows(new_ids, ids, mask)

    def generate(self, prompt: str, max_new_tokens: int = 512, temperature: float = 0.0, **generation_kwargs) -> str:
        return self.generate_batch([prompt], max_new_tokens, temperature, **generation_kwargs)[0]

    # -- teardown -----------------------------------------------------------------------------
    def cleanup(self):
        self.hooks = []
        for name in ("engine", "model", "tokenizer", "last_session"):
            if hasattr(self, name):
                delattr(self, name)
        gc.collect()
        if torch.cuda.is_available():
            torch.cuda.empty_cache()


def get_layer_at_fraction(model_wrapper, fraction: float) -> int:
    """`int(n_layers * fraction)` clamped to a valid index (reference model_utils.py:903-916)."""
    return max(0, min(int(model_wrapper.n_layers * fraction), model_wrapper.n_layers - 1))


def load_model(model_name: str, device: str = "cuda", dtype: str = "bfloat16", quantization: Optional[str] = None
               ) -> ModelWrapper:
    """reference model_utils.py:919-966; `arch:<name>` builds a random-init model offline."""
    torch_dtype = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}.get(dtype, torch.bfloat16)
    if quantization is not None:
        raise NotImplementedError("bitsandbytes quantisation is outside the B200 hot path")
    if model_name.startswith("arch:"):
        return ModelWrapper.from_arch(model_name[5:], device=device, dtype=torch_dtype)
    return ModelWrapper(model_name=model_name, device=device, dtype=torch_dtype)
