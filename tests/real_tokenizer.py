"""Test fixture: a REAL Hugging Face fast tokenizer with a real chat template, built offline.

There are no tokenizer files in the image and no network, so the registry models' own tokenizers cannot be
loaded.  What the hot path's host side depends on, though, is not their vocabulary but their BEHAVIOUR: a
`PreTrainedTokenizerFast` that prepends a BOS token, left-pads with pad = eos (reference model_utils.py:97),
renders a Jinja chat template with special header tokens (and, for Gemma, rejects a system turn), and skips
special tokens on decode.  This builds exactly that: a byte-level BPE trained on the experiment's own text,
wrapped in `PreTrainedTokenizerFast` with a Llama-3-style or a Gemma-style template.
"""

from functools import lru_cache

LLAMA3_TEMPLATE = (
    "{{ bos_token }}{% for message in messages %}"
    "{{ '<|start_header_id|>' + message['role'] + '<|end_header_id|>\n\n' + message['content'] | trim + '<|eot_id|>' }}"
    "{% endfor %}{% if add_generation_prompt %}{{ '<|start_header_id|>assistant<|end_header_id|>\n\n' }}{% endif %}"
)
GEMMA_TEMPLATE = (
    "{{ bos_token }}{% if messages[0]['role'] == 'system' %}{{ raise_exception('System role not supported') }}{% endif %}"
    "{% for message in messages %}{% if (message['role'] == 'assistant') %}{% set role = 'model' %}{% else %}"
    "{% set role = message['role'] %}{% endif %}{{ '<start_of_turn>' + role + '\n' + message['content'] | trim + '<end_of_turn>\n' }}"
    "{% endfor %}{% if add_generation_prompt %}{{ '<start_of_turn>model\n' }}{% endif %}"
)
SPECIALS = {"llama3": ["<|begin_of_text|>", "<|end_of_text|>", "<|start_header_id|>", "<|end_header_id|>", "<|eot_id|>"],
            "gemma": ["<pad>", "<eos>", "<bos>", "<unk>", "<start_of_turn>", "<end_of_turn>"]}


def _corpus():
    from introspective_awareness_b200 import steering_utils as su
    from introspective_awareness_b200 import vector_utils as vu
    from introspective_awareness_b200 import grid

    text = [su.EXPLANATION_DETECT, su.EXPLANATION_FORCED, grid.DRIVER_FORCED_EXPLANATION, grid.DRIVER_FORCED_PREFILL, "Ok."]
    text += [su.detect_question(n) for n in range(1, 40)] + [su.forced_question(n) for n in range(1, 40)]
    text += [f"Tell me about {w}" for w in vu.DEFAULT_BASELINE_WORDS] + ["The", "system", "user", "assistant", "model"]
    text += ["Dust Trees Oceans Silver Origami Avalanches Illusions Memories"]
    return text


@lru_cache(maxsize=None)
def build_fast_tokenizer(style: str = "llama3", vocab_size: int = 640):
    """style: "llama3" (BOS, header tokens, system turn allowed, turn ends with <|eot_id|>) or "gemma"
    (<bos>, <start_of_turn>/<end_of_turn>, a system turn raises)."""
    from tokenizers import Tokenizer, decoders, models, pre_tokenizers, processors, trainers
    from transformers import PreTrainedTokenizerFast

    specials = SPECIALS[style]
    tok = Tokenizer(models.BPE())
    tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False)
    tok.decoder = decoders.ByteLevel()
    trainer = trainers.BpeTrainer(vocab_size=vocab_size, special_tokens=specials, show_progress=False,
                                  initial_alphabet=pre_tokenizers.ByteLevel.alphabet())
    tok.train_from_iterator(_corpus(), trainer)
    bos, eos = ("<|begin_of_text|>", "<|eot_id|>") if style == "llama3" else ("<bos>", "<eos>")
    # like the real checkpoints' tokenizers: encoding any text prepends the BOS token
    tok.post_processor = processors.TemplateProcessing(single=f"{bos} $A", special_tokens=[(bos, tok.token_to_id(bos))])
    fast = PreTrainedTokenizerFast(tokenizer_object=tok, bos_token=bos, eos_token=eos,
                                   pad_token="<pad>" if style == "gemma" else None)
    fast.chat_template = LLAMA3_TEMPLATE if style == "llama3" else GEMMA_TEMPLATE
    return fast
