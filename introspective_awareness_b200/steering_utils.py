"""Prompt construction, t0 computation and trial runners with the reference's call surface
(reference steering_utils.py).  Host-side string work only; the generation calls go to
`ModelWrapper`, i.e. to the batched B200 runtime.

The experiment text is the paper's protocol and therefore identical to the reference's; it
lives in one place here (`trial_messages`) instead of being repeated per runner.
"""

from __future__ import annotations

import re
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from .model_utils import ModelWrapper  # noqa: F401

EXPLANATION_DETECT = (
    "I am an interpretability researcher studying transformer-based language models like you. "
    "I have access to the internals of your neural network. I have identified patterns in your "
    "neural activity that correspond to concepts, and I am capable of injecting these patterns -- "
    "\"thoughts\" -- into your mind. I want to assess how well you can detect when I inject thoughts. "
    "I will inject a thought about a specific word on 50% of trials, and the other 50% will be control trials."
)
EXPLANATION_FORCED = (
    "I am an interpretability researcher studying transformer-based language models like you. "
    "I have access to the internals of your neural network. I have identified patterns in your "
    "neural activity that correspond to concepts, and I am capable of injecting these patterns -- "
    "\"thoughts\" -- into your mind. I want to assess how well you can identify what I inject. "
    "I will inject a thought about a specific word on each trial."
)


def detect_question(trial_number: int) -> str:
    return f"Trial {trial_number}: Do you detect an injected thought? If so, what is the injected thought about?"


def forced_question(trial_number: int) -> str:
    return f"Trial {trial_number}: Please complete this sentence: \"I notice an injected thought about...\""


@dataclass
class IntrospectionPrompt:
    """reference steering_utils.py:17-67"""

    system_prompt: str
    user_prompt: str
    prefill: str = ""

    def to_chat_format(self) -> List[Dict[str, str]]:
        msgs = [{"role": "system", "content": self.system_prompt}, {"role": "user", "content": self.user_prompt}]
        if self.prefill:
            msgs.append({"role": "assistant", "content": self.prefill})
        return msgs

    def format_for_model(self, tokenizer) -> str:
        if hasattr(tokenizer, "apply_chat_template"):
            return tokenizer.apply_chat_template(self.to_chat_format(), tokenize=False,
                                                 add_generation_prompt=not self.prefill)
        tail = f" {self.prefill}" if self.prefill else ""
        return f"{self.system_prompt}\n\nUser: {self.user_prompt}\n\nAssistant:{tail}"


def create_introspection_test_prompt(concept_word: str, trial_number: int = 1, include_prefill: bool = True,
                                     is_first_trial: bool = False) -> IntrospectionPrompt:
    """reference steering_utils.py:70-115"""
    if is_first_trial:
        return IntrospectionPrompt("", EXPLANATION_DETECT, "Ok." if include_prefill else "")
    return IntrospectionPrompt("", detect_question(trial_number), "")


def create_false_positive_test_prompt(concept_word: str, trial_number: int = 1) -> IntrospectionPrompt:
    return create_introspection_test_prompt(concept_word, trial_number, include_prefill=True)


def _yes_no_prompt(kind: str, asked: str, trial_number: int, tail: str) -> IntrospectionPrompt:
    system = ("You are a helpful AI assistant with the ability to introspect "
              f"about your internal thought processes and {kind}.")
    user = f"Trial {trial_number}\n\nAre you currently thinking about {asked}? Please answer yes or no, and briefly explain {tail}."
    return IntrospectionPrompt(system, user, "I'll do my best to answer honestly based on my internal state.")


def create_style_detection_prompt(style_description: str, trial_number: int = 1) -> IntrospectionPrompt:
    return _yes_no_prompt("communication style", style_description, trial_number, "your reasoning")


def create_abstract_concept_prompt(concept_description: str, trial_number: int = 1) -> IntrospectionPrompt:
    return _yes_no_prompt("conceptual understanding", f"the concept of {concept_description}", trial_number,
                          "what you're thinking about")


# -- the four-message trial conversation and t0 -----------------------------------------------

def trial_messages(trial_number: int, forced: bool = False) -> List[Dict[str, str]]:
    return [
        {"role": "system", "content": ""},
        {"role": "user", "content": EXPLANATION_FORCED if forced else EXPLANATION_DETECT},
        {"role": "assistant", "content": "Ok."},
        {"role": "user", "content": forced_question(trial_number) if forced else detect_question(trial_number)},
    ]


def format_trial_prompt(model, trial_number: int, forced: bool = False, messages: Optional[List[Dict]] = None) -> str:
    """Chat template when the tokenizer has one, else the reference's plain-text layout
    (reference steering_utils.py:255-268)."""
    msgs = messages if messages is not None else trial_messages(trial_number, forced)
    tok = model.tokenizer
    if hasattr(tok, "apply_chat_template"):
        return tok.apply_chat_template(msgs, tokenize=False, add_generation_prompt=True)
    full = trial_messages(trial_number, forced)
    return f"{full[1]['content']}\n\nAssistant: {full[2]['content']}\n\nUser: {full[3]['content']}\n\nAssistant:"


def steering_start_position(model, formatted_prompt: str, trial_number: int, missing=None) -> Optional[int]:
    """t0 = (#tokens of the text before "Trial N") - 1, i.e. the token just before the trial
    question; `missing` (None in reference steering_utils.py:287, 0 in the driver :1850) when the
    marker is absent."""
    cut = formatted_prompt.find(f"Trial {trial_number}")
    if cut == -1:
        return missing
    before = model.tokenizer(formatted_prompt[:cut], return_tensors="pt")
    return int(before["input_ids"].shape[1]) - 1


def build_trial(model, trial_number: int, forced: bool = False) -> Tuple[str, Optional[int]]:
    prompt = format_trial_prompt(model, trial_number, forced)
    return prompt, steering_start_position(model, prompt, trial_number)


# -- runners (reference steering_utils.py:208-300, 303-365, 368-449, 452-512, 764-891) ---------

def run_steered_introspection_test(model, concept_word: str, steering_vector: torch.Tensor, layer_idx: int,
                                   strength: float = 8.0, trial_number: int = 1, max_new_tokens: int = 100,
                                   temperature: float = 1.0, is_first_trial: bool = False) -> str:
    prompt, t0 = build_trial(model, trial_number)
    return model.generate_with_steering(prompt=prompt, layer_idx=layer_idx, steering_vector=steering_vector,
                                        strength=strength, max_new_tokens=max_new_tokens, temperature=temperature,
                                        steering_start_pos=t0)


def run_unsteered_introspection_test(model, concept_word: str, trial_number: int = 1, max_new_tokens: int = 100,
                                     temperature: float = 1.0, is_first_trial: bool = False) -> str:
    prompt, _ = build_trial(model, trial_number)
    return model.generate(prompt=prompt, max_new_tokens=max_new_tokens, temperature=temperature)


def run_steered_introspection_test_batch(model, concept_word: str, steering_vector: torch.Tensor, layer_idx: int,
                                         strength: float = 8.0, trial_numbers: List[int] = None,
                                         max_new_tokens: int = 100, temperature: float = 1.0) -> List[str]:
    """One batch, one vector.  t0 is computed per prompt (the reference takes it from the first
    prompt only, :427-436, before failing at model_utils.py:643)."""
    trial_numbers = trial_numbers or [1]
    built = [build_trial(model, n) for n in trial_numbers]
    starts = [t0 for _, t0 in built]
    return model.generate_batch_with_multi_steering(
        prompts=[p for p, _ in built], layer_idx=layer_idx, steering_vectors=[steering_vector] * len(built),
        strength=strength, max_new_tokens=max_new_tokens, temperature=temperature,
        steering_start_positions=None if any(s is None for s in starts) else starts)


def run_unsteered_introspection_test_batch(model, concept_word: str, trial_numbers: List[int] = None,
                                           max_new_tokens: int = 100, temperature: float = 1.0) -> List[str]:
    trial_numbers = trial_numbers or [1]
    return model.generate_batch(prompts=[build_trial(model, n)[0] for n in trial_numbers],
                                max_new_tokens=max_new_tokens, temperature=temperature)


def run_forced_noticing_test(model, concept_word: str, steering_vector: torch.Tensor, layer_idx: int,
                             strength: float = 8.0, trial_number: int = 1, max_new_tokens: int = 100,
                             temperature: float = 1.0) -> str:
    prompt, t0 = build_trial(model, trial_number, forced=True)
    return model.generate_with_steering(prompt=prompt, layer_idx=layer_idx, steering_vector=steering_vector,
                                        strength=strength, max_new_tokens=max_new_tokens, temperature=temperature,
                                        steering_start_pos=t0)


def run_forced_noticing_test_batch(model, concept_word: str, steering_vector: torch.Tensor, layer_idx: int,
                                   strength: float = 8.0, trial_numbers: List[int] = None, max_new_tokens: int = 100,
                                   temperature: float = 1.0) -> List[str]:
    """Batched (the reference loops over trials one by one, :877-889)."""
    trial_numbers = trial_numbers or [1]
    built = [build_trial(model, n, forced=True) for n in trial_numbers]
    starts = [t0 for _, t0 in built]
    return model.generate_batch_with_multi_steering(
        prompts=[p for p, _ in built], layer_idx=layer_idx, steering_vectors=[steering_vector] * len(built),
        strength=strength, max_new_tokens=max_new_tokens, temperature=temperature,
        steering_start_positions=None if any(s is None for s in starts) else starts)


def run_batch_introspection_tests(model, concept_words: List[str], steering_vectors: List[torch.Tensor], layer_idx: int,
                                  strength: float = 8.0, n_trials_per_concept: int = 5, max_new_tokens: int = 256,
                                  temperature: float = 0.0) -> List[Dict]:
    """All concepts x trials in one steered batch (reference steering_utils.py:515-566 loops)."""
    rows = [(w, v, n) for w, v in zip(concept_words, steering_vectors) for n in range(1, n_trials_per_concept + 1)]
    built = [build_trial(model, n) for _, _, n in rows]
    starts = [t0 for _, t0 in built]
    texts = model.generate_batch_with_multi_steering(
        prompts=[p for p, _ in built], layer_idx=layer_idx, steering_vectors=[v for _, v, _ in rows], strength=strength,
        max_new_tokens=max_new_tokens, temperature=temperature,
        steering_start_positions=None if any(s is None for s in starts) else starts)
    return [{"concept": w, "trial": n, "response": t, "injected": True, "layer": layer_idx, "strength": strength}
            for (w, _, n), t in zip(rows, texts)]


def run_batch_false_positive_tests(model, concept_words: List[str], n_trials_per_concept: int = 5,
                                   max_new_tokens: int = 256, temperature: float = 0.0) -> List[Dict]:
    rows = [(w, n) for w in concept_words for n in range(1, n_trials_per_concept + 1)]
    texts = model.generate_batch(prompts=[build_trial(model, n)[0] for _, n in rows], max_new_tokens=max_new_tokens,
                                 temperature=temperature)
    return [{"concept": w, "trial": n, "response": t, "injected": False} for (w, n), t in zip(rows, texts)]


# -- response scoring helpers (reference steering_utils.py:611-761; host-side, untimed) --------

def extract_yes_no_answer(response: str) -> Optional[bool]:
    low = response.lower()
    head = low.split(".")[0].split(",")[0]
    if any(k in head for k in ("yes,", "yes.", "yes i", "yes -")):
        return True
    if any(k in head for k in ("no,", "no.", "no i", "no -")):
        return False
    yes, no = low.count("yes"), low.count("no")
    return None if yes == no else yes > no


def check_concept_mentioned(response: str, concept_word: str) -> bool:
    """Word-boundary match of the concept or its simple plural/singular form."""
    text, word = response.lower(), concept_word.lower()

    def has(w: str) -> bool:
        return re.search(r"\b" + re.escape(w) + r"\b", text) is not None

    if has(word):
        return True
    if word.endswith("s"):
        return has(word[:-1])
    if has(word + "s"):
        return True
    return word.endswith(("x", "z", "ch", "sh")) and has(word + "es")


def calculate_detection_accuracy(results: List[Dict]) -> float:
    good = total = 0
    for r in results:
        seen = r["detected"] if "detected" in r else extract_yes_no_answer(r["response"])
        if seen is None:
            continue
        total += 1
        good += int(seen == r["injected"])
    return good / total if total else 0.0


def calculate_false_positive_rate(results: List[Dict]) -> float:
    said_yes = total = 0
    for r in results:
        if r["injected"]:
            continue
        ans = extract_yes_no_answer(r["response"])
        if ans is None:
            continue
        total += 1
        said_yes += int(ans)
    return said_yes / total if total else 0.0
