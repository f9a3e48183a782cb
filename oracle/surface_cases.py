"""ORACLE — test infrastructure.  The calls `oracle/make_surface_golden.py` makes on the real reference and
`tests/test_surface_gpu.py` repeats on this package (pure data: importable without /root/reference)."""

# (tag, kwargs); a ("BASELINE", n) value stands for get_baseline_words(n) of whichever module is under test
CALLS = [
    ("posneg", dict(positive_prompts=["HI! HOW ARE YOU?", "I AM SHOUTING NOW", "LOUD WORDS"],
                    negative_prompts=["Hi! How are you?", "I am speaking now"], normalize=False)),
    ("posneg_norm", dict(positive_prompts=["HI! HOW ARE YOU?", "I AM SHOUTING NOW"],
                         negative_prompts=["Hi! How are you?", "I am speaking now", "quiet words"], normalize=True)),
    ("with_baseline", dict(concept_word="Dust", baseline_words=("BASELINE", 12), normalize=False)),
    ("with_baseline_tmpl", dict(concept_word="Oceans", baseline_words=("BASELINE", 5),
                                template="Think about {word} now", token_idx=-2, normalize=True)),
    ("simple", dict(concept_word="Trees", normalize=False)),
    ("simple_ctrl", dict(concept_word="Trees", control_prompt="A thing", template="the word {word}", normalize=True)),
    ("no_baseline", dict(concept_word="Silver", normalize=False)),
    ("no_baseline_norm", dict(concept_word="Silver", template="Say {word}", normalize=True)),
]
FUNCS = {"posneg": "extract_concept_vector", "with_baseline": "extract_concept_vector_with_baseline",
         "simple": "extract_concept_vector_simple", "no_baseline": "extract_concept_vector_no_baseline"}
UNDERSPEC = dict(target="Dust", related=["Trees", "Oceans", "Dust"], n_baseline=8)


def func_of(tag: str) -> str:
    return FUNCS[max((k for k in FUNCS if tag.startswith(k)), key=len)]


def materialize(kwargs: dict, get_baseline_words) -> dict:
    return {k: (get_baseline_words(v[1]) if isinstance(v, tuple) and v and v[0] == "BASELINE" else v) for k, v in kwargs.items()}
