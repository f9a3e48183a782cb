"""GPU checks of the runtime-support kernels that are not hot-path rows: the fused rotary
embedding must produce the bits of transformers' eager apply_rotary_pos_emb."""

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("B,T,nh,nkv,D", [(200, 1, 32, 8, 128), (3, 37, 14, 2, 64), (2, 16, 16, 8, 256), (1, 5, 4, 4, 32)])
def test_fused_rope_is_bit_identical_to_hf(dtype, B, T, nh, nkv, D):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from transformers.models.llama.modeling_llama import apply_rotary_pos_emb

    from introspective_awareness_b200 import kernels

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(B * 1000 + T)
    q = torch.randn(B, T, nh * D, device=dev, generator=g).to(dtype).view(B, T, nh, D).transpose(1, 2)
    k = torch.randn(B, T, nkv * D, device=dev, generator=g).to(dtype).view(B, T, nkv, D).transpose(1, 2)
    ang = torch.rand(B, T, D // 2, device=dev, generator=g) * 6.28
    emb = torch.cat((ang, ang), dim=-1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    want_q, want_k = apply_rotary_pos_emb(q, k, cos, sin)
    got = kernels.rope(q, k, cos, sin)
    assert got is not None
    torch.cuda.synchronize()
    assert got[0].stride() == q.stride()
    assert torch.equal(got[0], want_q) and torch.equal(got[1], want_k)
    # a layout the kernel does not cover is declined, not mis-read
    if T > 1 and nh > 1:
        assert kernels.rope(q.contiguous(), k, cos, sin) is None


def test_graph_capture_survives_a_garbage_collection_of_older_sessions(monkeypatch):
    """Old sessions die through the cyclic collector, whenever it happens to run; finalising their CUDA
    graphs inside a new capture invalidates it.  Arrange exactly that: let a used wrapper become
    garbage while collection is held off, then make the collector eager from the moment the next
    capture begins.  The engine must have collected beforehand and must keep the collector out."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    import gc

    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids

    def run_once(rows):
        w = ModelWrapper.from_arch("tiny-llama", device="cuda:0", dtype=torch.bfloat16, seed=2)
        ids, mask, _ = synthetic_prompt_ids(rows, 12, 6, 512, seed=3)
        table = torch.randn(2, 128, generator=torch.Generator().manual_seed(0)).to(torch.bfloat16)
        return w.steered_generate_ids(ids, mask, 1, table, [0, 1] * (rows // 2), [4.0] * rows, [0] * rows, 6).cpu().clone()

    old = gc.get_threshold()
    real_begin = torch.cuda.CUDAGraph.capture_begin

    def eager_collector_begin(self, *a, **k):
        real_begin(self, *a, **k)
        gc.set_threshold(1, 1, 1)

    try:
        gc.collect()
        gc.set_threshold(10 ** 9, 10 ** 9, 10 ** 9)   # nothing is collected on its own from here on
        first = run_once(4)                           # wrapper, engine, session and 5 graphs are now garbage
        monkeypatch.setattr(torch.cuda.CUDAGraph, "capture_begin", eager_collector_begin)
        second = run_once(6)
        assert first.shape == (4, 6) and second.shape == (6, 6)
    finally:
        gc.set_threshold(*old)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("B,T,nh,nkv,D,pos0,fused", [(200, 1, 32, 8, 128, 140, True), (3, 37, 14, 2, 64, 0, False),
                                                     (2, 16, 16, 8, 256, 5, True), (5, 1, 4, 4, 32, 0, False)])
def test_rope_append_matches_hf_rotary_plus_cache_update(dtype, B, T, nh, nkv, D, pos0, fused):
    """One launch = apply_rotary_pos_emb + the two copies of `past_key_values.update`, bit for bit; inputs
    either three dense projections or column slices of one fused projection output."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from transformers.models.llama.modeling_llama import apply_rotary_pos_emb

    from introspective_awareness_b200 import kernels

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(B * 100 + T)
    n = (nh + 2 * nkv) * D
    qkv = torch.randn(B, T, n, device=dev, generator=g).to(dtype)
    q, k, v = qkv.split([nh * D, nkv * D, nkv * D], dim=-1)
    if not fused:
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    ang = torch.rand(B, T, D // 2, device=dev, generator=g) * 6.28
    emb = torch.cat((ang, ang), dim=-1)
    cos, sin = emb.cos().to(dtype), emb.sin().to(dtype)
    L = pos0 + T + 3
    kc = torch.full((B, nkv, L, D), 7.0, device=dev, dtype=dtype)
    vc = torch.full((B, nkv, L, D), 7.0, device=dev, dtype=dtype)
    want_q, want_k = apply_rotary_pos_emb(q.reshape(B, T, nh, D).transpose(1, 2), k.reshape(B, T, nkv, D).transpose(1, 2), cos, sin)
    want_kc, want_vc = kc.clone(), vc.clone()
    want_kc[:, :, pos0:pos0 + T] = want_k
    want_vc[:, :, pos0:pos0 + T] = v.reshape(B, T, nkv, D).transpose(1, 2)
    got_q = kernels.rope_append(q, k, v, cos, sin, kc, vc, pos0, nh, nkv, D)
    torch.cuda.synchronize()
    assert got_q.shape == want_q.shape and torch.equal(got_q, want_q)
    assert torch.equal(kc, want_kc) and torch.equal(vc, want_vc)   # slot written, everything else untouched
    # cache-less form: destination is a transposed view of a dense [B, T, n_kv, D] buffer
    kd = torch.empty(B, T, nkv, D, device=dev, dtype=dtype).transpose(1, 2)
    vd = torch.empty(B, T, nkv, D, device=dev, dtype=dtype).transpose(1, 2)
    kernels.rope_append(q, k, v, cos, sin, kd, vd, 0, nh, nkv, D)
    assert torch.equal(kd, want_k) and torch.equal(vd, v.reshape(B, T, nkv, D).transpose(1, 2))
    with pytest.raises(ValueError):
        kernels.rope_append(q, k, v, cos, sin, kc, vc, L - T + 1, nh, nkv, D)   # would run past the cache


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("rows,width,fused", [(200, 14336, True), (7, 4864, False), (3, 256, True), (25600, 1024, False),
                                              (8192, 8192, True), (5000, 14336, False)])  # the last two: the table kernel (>= 64 M elements)
def test_silu_mul_is_bit_identical_to_the_two_eager_kernels(dtype, rows, width, fused):
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import kernels

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(rows + width)
    gu = (torch.randn(rows, 2 * width, device=dev, generator=g) * 4).to(dtype)
    gu[0, :8] = torch.tensor([0.0, -0.0, 30.0, -30.0, 88.0, -88.0, 1e-3, -1e-3], device=dev).to(dtype)  # tails of exp
    gate, up = gu.split(width, dim=-1)
    if not fused:
        gate, up = gate.contiguous(), up.contiguous()
    want = torch.nn.functional.silu(gate) * up
    got = kernels.silu_mul(gate, up)
    torch.cuda.synchronize()
    assert got.is_contiguous() and torch.equal(got, want)
    got3 = kernels.silu_mul(gate.reshape(1, rows, width) if not fused else gate.unflatten(0, (1, rows)),
                            up.reshape(1, rows, width) if not fused else up.unflatten(0, (1, rows)))
    assert torch.equal(got3[0], want)


@pytest.mark.parametrize("arch", ["tiny-llama", "tiny-gemma2"])
def test_one_set_of_decode_graphs_serves_every_layer_and_sessions_are_kept(arch):
    """The steered layer is data in the decode graphs (one row of vector indices per decoder layer, -1 except at
    the steered one), so a sweep over layers re-uses the graphs of its batch shape; the last few batch shapes
    keep their sessions.  Every layer's tokens and logits must still be the hook path's."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids
    from oracle import ref_hooks

    w = ModelWrapper.from_arch(arch, device="cuda:0", dtype=torch.bfloat16, seed=23)
    V = w.model.config.vocab_size
    rows = 12
    ids, mask, lens = synthetic_prompt_ids(rows, 30, 20, V, seed=4)
    table = (torch.randn(4, w.engine.hidden, generator=torch.Generator().manual_seed(2)) * 0.5).to(torch.bfloat16)
    t0 = (lens - 9).tolist()
    row_vec = [i % 5 - 1 for i in range(rows)]  # includes control rows
    graphs_seen = None
    for layer in (0, 1, 2, 3, 1):
        ours = w.steered_generate_ids(ids, mask, layer, table, row_vec, [4.0] * rows, t0, 10, keep_logits=True).cpu().clone()
        logits = torch.stack(w.last_session.step_logits).cpu()
        sess = w.last_session
        assert len(w.engine._sessions) == 1 and len(sess.graphs) == 1            # one shape, one set of step graphs
        these = [id(g) for g in next(iter(sess.graphs.values()))]
        assert graphs_seen is None or these == graphs_seen, "graphs were captured again for another layer"
        graphs_seen = these
        vec_rows = [table[j] if j >= 0 else torch.zeros_like(table[0]) for j in row_vec]  # a zero vector = a control row
        ref = ref_hooks.steered_generate(w.model, ids, mask, layer, vec_rows, 4.0, t0, 10, want_logits=True)
        assert torch.equal(ours, ref["new_ids"]), layer
        assert (logits - ref["logits"]).abs().max().item() <= 2e-2, layer
    # other shapes get their own sessions; the least recently used one goes when max_sessions is exceeded
    first = w.last_session
    for width in (29, 28):
        w.steered_generate_ids(ids[:, -width:], mask[:, -width:], 1, table, row_vec, [4.0] * rows,
                               [max(0, t - (30 - width)) for t in t0], 10)
    assert len(w.engine._sessions) == 3 and first in w.engine._sessions.values()
    w.steered_generate_ids(ids, mask, 1, table, row_vec, [4.0] * rows, t0, 10)              # touch the first shape
    w.steered_generate_ids(ids[:, -27:], mask[:, -27:], 1, table, row_vec, [4.0] * rows, [max(0, t - 3) for t in t0], 10)
    assert len(w.engine._sessions) == 3 and first in w.engine._sessions.values()          # the width-29 session went
    assert [k[1] for k in w.engine._sessions] == [28, 30, 27]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("rows,width,fused", [(200, 14336, True), (7, 4864, False), (3, 256, True), (6000, 14336, False)])
def test_table_activation_gate_multiply_is_bit_identical_to_the_eager_kernels(dtype, rows, width, fused):
    """ia_act_mul_table with the table filled by ATen's own tanh-GELU (Gemma-2) and SiLU: act_fn(gate) * up, bit for
    bit, small launches (few CTAs) and large ones, dense operands or the two halves of one fused projection output."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import kernels

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(rows + width)
    gu = (torch.randn(rows, 2 * width, device=dev, generator=g) * 4).to(dtype)
    gu[0, :8] = torch.tensor([0.0, -0.0, 30.0, -30.0, 88.0, -88.0, 1e-3, -1e-3], device=dev).to(dtype)
    gu[0, 8:11] = torch.tensor([float("inf"), float("-inf"), float("nan")], device=dev).to(dtype)
    gate, up = gu.split(width, dim=-1)
    if not fused:
        gate, up = gate.contiguous(), up.contiguous()
    for act in (torch.nn.GELU(approximate="tanh"), torch.nn.SiLU()):
        table = kernels.activation_table(act, dtype, dev)
        assert table.shape == (65536,) and table.dtype == dtype
        want = act(gate) * up
        got = kernels.act_mul_table(gate, up, table)
        torch.cuda.synchronize()
        assert got.is_contiguous()
        nan = want.isnan()
        assert torch.equal(got.isnan(), nan), type(act).__name__
        assert torch.equal(got.view(torch.int16)[~nan], want.view(torch.int16)[~nan]), type(act).__name__  # bit patterns


def test_generation_config_settings_inside_the_decode_graphs():
    """Repetition penalty (a Qwen2.5-Instruct default, applied to greedy decoding too) and an EOS *set* whose second id
    fires, on the GPU where the decode steps are CUDA graphs: same tokens as HF `generate` with the same
    generation_config, for steered rows as for plain ones (the penalty buffer and the EOS ids live in the graphs)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200.model_utils import ModelWrapper
    from introspective_awareness_b200.synthetic import synthetic_prompt_ids
    from oracle import ref_hooks

    w = ModelWrapper.from_arch("tiny-llama", device="cuda:0", dtype=torch.bfloat16, seed=41)
    V = w.model.config.vocab_size
    ids, mask, lens = synthetic_prompt_ids(10, 24, 14, V, seed=6)
    gc_ = w.model.generation_config
    gc_.repetition_penalty = 1.3
    with torch.no_grad():
        free = w.model.generate(input_ids=ids.cuda(), attention_mask=mask.cuda(), max_new_tokens=12, do_sample=False, pad_token_id=0)
    fires = int(free[0, 24 + 4])  # row 0's fifth new token becomes an end-of-sequence id
    gc_.eos_token_id = [V - 1, fires]
    try:
        with torch.no_grad():
            ref = w.model.generate(input_ids=ids.cuda(), attention_mask=mask.cuda(), max_new_tokens=12, do_sample=False, pad_token_id=0)
        got = w.steered_generate_ids(ids, mask, None, None, None, None, None, 12)
        new = ref[:, 24:]
        assert torch.equal(got[:, : new.shape[1]], new) and (got[:, new.shape[1]:] == 0).all()
        assert int(got[0, 4]) == fires and (got[0, 5:] == 0).all()
        # steered rows under the same settings: the hook path through HF generate
        table = (torch.randn(3, w.engine.hidden, generator=torch.Generator().manual_seed(1)) * 0.5).to(torch.bfloat16)
        row_vec = [i % 3 for i in range(10)]
        t0 = (lens - 6).tolist()
        want = ref_hooks.steered_generate(w.model, ids, mask, 1, [table[j] for j in row_vec], 4.0, t0, 12)["new_ids"]
        mine = w.steered_generate_ids(ids, mask, 1, table, row_vec, [4.0] * 10, t0, 12).cpu()
        assert torch.equal(mine[:, : want.shape[1]], want) and (mine[:, want.shape[1]:] == 0).all()
    finally:
        gc_.repetition_penalty, gc_.eos_token_id = 1.0, None
