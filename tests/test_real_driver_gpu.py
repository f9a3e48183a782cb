"""north_star: "... so that detect_injected_thoughts.py and model_utils.py drive it unchanged."

The reference's sweep driver, UNMODIFIED (compiled into oracle/_ref by oracle/build_ref.py), is run twice on the GPU
in sweep mode (`main()`, --no-llm-judge, temperature 0): once bound to the reference's own model_utils / vector_utils /
steering_utils with the reference's ModelWrapper around a seeded model, once with those three names pointed at this
package (INTEGRATION.md level 1, the import swap) and this package's ModelWrapper around the same weights.  Everything
the driver writes must come out the same: every result row of every (layer, strength) configuration — responses
included, i.e. every generated token — the metrics, the CSVs and the saved concept vectors.
"""

import json
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(drv, make_model, out_dir, model_name, concepts, fractions, strengths):
    drv.load_model = lambda **kw: make_model()
    argv = sys.argv
    sys.argv = ["detect_injected_thoughts.py", "-m", model_name, "-c", *concepts, "-nb", "8", "-ls", *map(str, fractions),
                "-ss", *map(str, strengths), "-nt", "5", "-t", "0.0", "-mt", "12", "-bs", "4", "-od", str(out_dir), "-nlj", "-ow",
                "-d", "cuda", "-dt", "bfloat16"]
    try:
        drv.main()
    finally:
        sys.argv = argv


@pytest.mark.parametrize("arch,chat", [("tiny-llama", False), ("tiny-qwen2", True), ("tiny-gemma2", True)])
def test_the_unmodified_driver_gives_the_same_files_on_this_package(tmp_path, arch, chat):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import introspective_awareness_b200 as ia
    from introspective_awareness_b200.stub_tokenizer import StubTokenizer
    from introspective_awareness_b200.synthetic import build_model
    from oracle import build_ref, real_ref

    if not os.path.exists(os.path.join(real_ref.REF_DIR, "detect_injected_thoughts.bin")):
        build_ref.build(verbose=False)
    mods = real_ref.load()
    if mods is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    mu, vu = mods
    ref_drv = real_ref.load_driver({"model_utils": mu, "vector_utils": vu, "steering_utils": real_ref.load_steering_utils()})
    our_drv = real_ref.load_driver({"model_utils": ia.model_utils, "vector_utils": ia.vector_utils, "steering_utils": ia.steering_utils})
    assert ref_drv is not None and our_drv is not None
    assert our_drv.extract_concept_vectors_batch is ia.vector_utils.extract_concept_vectors_batch
    assert ref_drv.extract_concept_vectors_batch is vu.extract_concept_vectors_batch

    name = "gemma2_9b" if "gemma" in arch else "tiny_model"  # the registry name switches the driver's system-turn filter
    concepts, fractions, strengths = ["Dust", "Trees", "Oceans"], [0.3, 0.8], [2.0, 8.0]

    def weights():
        return build_model(arch, dtype=torch.bfloat16, device="cuda:0", seed=77)

    def tok(model):
        return StubTokenizer(vocab_size=model.config.vocab_size, chat_template=chat)

    def reference_model():
        m = weights()
        return real_ref.wrap(m, "cuda:0", torch.bfloat16, tokenizer=tok(m), model_name=name)

    def our_model():
        m = weights()
        return ia.model_utils.ModelWrapper(name, device="cuda:0", dtype=torch.bfloat16, model=m, tokenizer=tok(m), compat_slice=True)

    _run(ref_drv, reference_model, tmp_path / "ref", name, concepts, fractions, strengths)
    _run(our_drv, our_model, tmp_path / "ours", name, concepts, fractions, strengths)

    n_rows = 0
    for f in fractions:
        for a in strengths:
            cfg = f"layer_{f:.2f}_strength_{a}"
            theirs = json.load(open(tmp_path / "ref" / name / cfg / "results.json"))
            ours = json.load(open(tmp_path / "ours" / name / cfg / "results.json"))
            assert ours["results"] == theirs["results"], cfg   # every row: prompts' trial numbers, responses, detected flags
            assert ours["metrics"] == theirs["metrics"], cfg
            assert open(tmp_path / "ours" / name / cfg / "results.csv").read() == open(tmp_path / "ref" / name / cfg / "results.csv").read()
            n_rows += len(ours["results"])
            assert {r["trial_type"] for r in ours["results"]} == {"injection", "control", "forced_injection"}
            assert all(r["response"] for r in ours["results"])
        for c in concepts:
            a_, b_ = (torch.load(tmp_path / side / name / "vectors" / f"layer_{f:.2f}" / f"{c}.pt") for side in ("ours", "ref"))
            assert a_.dtype == torch.bfloat16 and torch.equal(a_, b_), (f, c)
    assert n_rows == len(fractions) * len(strengths) * len(concepts) * (2 + 3 + 2)
