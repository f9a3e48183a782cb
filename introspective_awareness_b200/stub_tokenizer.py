"""Deterministic offline tokenizer used wherever no real tokenizer is on disk.

There is no network and no HF cache on the build or GPU boxes, so the reference's
`AutoTokenizer.from_pretrained` (reference model_utils.py:91-101) cannot run.  This
stand-in exposes exactly the attributes the reference's hot path touches:

* `__call__(str | list[str], return_tensors="pt", padding=..., truncation=...)`
  returning a mapping with `input_ids` / `attention_mask` and a `.to(device)`
  method (reference model_utils.py:329-334, 413, 816);
* left padding (`padding_side = "left"`, reference model_utils.py:97);
* `decode(ids, skip_special_tokens=True)` (reference model_utils.py:449-451, 874-876);
* `pad_token_id`, `eos_token_id`, `pad_token`, `eos_token`, `__len__`.

It deliberately has NO `apply_chat_template` attribute unless `chat_template=True`,
so the reference helpers take their plain-text fallback
(reference vector_utils.py:516-517, steering_utils.py:261-268).

Tokenisation is word-level: text is split into runs of word characters, single
punctuation marks and "\n\n" paragraph breaks; each piece maps to
`2 + crc32(piece) % (vocab_size - 2)`.  A piece-prefix therefore always tokenises to a
prefix of the full tokenisation, which is the property the reference's t0 computation
relies on (reference steering_utils.py:276-287).
"""

from __future__ import annotations

import re
import zlib
from typing import Dict, List, Sequence, Union

import torch

_PIECE = re.compile(r"\n\n|\w+|[^\w\s]")


class BatchEncoding(dict):
    """dict with `.to(device)`, like transformers.BatchEncoding."""

    def to(self, device):
        return BatchEncoding({k: v.to(device) for k, v in self.items()})

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError as exc:  # pragma: no cover - attribute protocol
            raise AttributeError(name) from exc


class StubTokenizer:
    pad_token = "<pad>"
    eos_token = "<eos>"
    pad_token_id = 0
    eos_token_id = 1

    def __init__(self, vocab_size: int = 512, chat_template: bool = False, max_length: int = 4096):
        if vocab_size < 8:
            raise ValueError("vocab_size must be >= 8")
        self.vocab_size = int(vocab_size)
        self.padding_side = "left"
        self.model_max_length = int(max_length)
        if chat_template:
            # Bound as an instance attribute so hasattr() is False for the default stub.
            self.apply_chat_template = self._apply_chat_template

    def __len__(self) -> int:
        return self.vocab_size

    # -- encoding ---------------------------------------------------------
    def encode(self, text: str) -> List[int]:
        return [2 + zlib.crc32(p.encode("utf-8")) % (self.vocab_size - 2) for p in _PIECE.findall(text)]

    def __call__(
        self,
        text: Union[str, Sequence[str]],
        return_tensors: str = "pt",
        padding: bool = False,
        truncation: bool = False,
        **_unused,
    ) -> BatchEncoding:
        single = isinstance(text, str)
        rows = [self.encode(text)] if single else [self.encode(t) for t in text]
        if truncation:
            rows = [r[: self.model_max_length] for r in rows]
        width = max((len(r) for r in rows), default=0)
        if not single and not padding and any(len(r) != width for r in rows):
            raise ValueError("ragged batch needs padding=True")
        ids = torch.full((len(rows), width), self.pad_token_id, dtype=torch.long)
        mask = torch.zeros((len(rows), width), dtype=torch.long)
        for i, r in enumerate(rows):
            if not r:
                continue
            if self.padding_side == "left":
                ids[i, width - len(r):] = torch.tensor(r, dtype=torch.long)
                mask[i, width - len(r):] = 1
            else:
                ids[i, : len(r)] = torch.tensor(r, dtype=torch.long)
                mask[i, : len(r)] = 1
        return BatchEncoding(input_ids=ids, attention_mask=mask)

    # -- decoding ---------------------------------------------------------
    def decode(self, ids, skip_special_tokens: bool = False) -> str:
        if isinstance(ids, torch.Tensor):
            ids = ids.tolist()
        out = []
        for t in ids:
            t = int(t)
            if skip_special_tokens and t in (self.pad_token_id, self.eos_token_id):
                continue
            out.append(f"t{t}")
        return " ".join(out)

    def batch_decode(self, batch, skip_special_tokens: bool = False) -> List[str]:
        return [self.decode(r, skip_special_tokens=skip_special_tokens) for r in batch]

    # -- optional chat template (only bound when chat_template=True) ------
    def _apply_chat_template(self, messages: List[Dict[str, str]], tokenize: bool = False,
                             add_generation_prompt: bool = True) -> str:
        parts = []
        for m in messages:
            if m.get("content", "") == "" and m.get("role") == "system":
                continue
            parts.append(f"<|{m['role']}|>\n\n{m['content']}\n\n")
        if add_generation_prompt:
            parts.append("<|assistant|>\n\n")
        text = "".join(parts)
        return self.encode(text) if tokenize else text
