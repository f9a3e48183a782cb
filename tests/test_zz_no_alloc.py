"""GPU test of ia_gather_meandiff_static, the K1 entry point for callers that cannot allocate.

Who that is: hosts that may touch only the device buffers handed to them — a plugin inside another
runtime that has already divided up GPU memory, code running where cudaMalloc's device-wide
synchronisation is forbidden, a bare FFI binding without an allocator.  They pass inputs and
outputs; the scratch is a fixed 4 MiB array inside the library.

The test drives the entry point through the C ABI with raw pointers and compares it with
ia_gather_meandiff (caller-supplied workspace) bit for bit — the tolerance every K1-vs-K1 check in
this suite uses — on two consecutive calls with different shapes, so that whatever the first call
left in the shared array would show up in the second.  It lives in its own file, collected last.
"""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(dev, L, C, Nb, H, T, seed, dtype=torch.bfloat16):
    g = torch.Generator(device=dev).manual_seed(seed)
    conc = [(torch.randn(C, T, H, device=dev, generator=g) * 3 + 5).to(dtype) for _ in range(L)]
    base = [(torch.randn(Nb, T, H, device=dev, generator=g) * 3 + 5).to(dtype) for _ in range(L)]
    return conc, base


def _run_static(lib, conc, base, mode, normalize, kernels, check_device_free=False):
    """All buffers are made here, by the caller; the call under test receives pointers only."""
    dev = conc[0].device
    L, (C, T, H) = len(conc), conc[0].shape
    Nb = base[0].shape[0]
    n_out = 1 if mode == "posneg" else C
    cptr, bptr = kernels._ptr_table(conc, "c"), kernels._ptr_table(base, "b")
    out32 = torch.full((L, n_out, H), float("nan"), dtype=torch.float32, device=dev)
    outr = torch.full((L, n_out, H), float("nan"), dtype=conc[0].dtype, device=dev)
    outn = torch.full((L, n_out), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    free_before, _ = torch.cuda.mem_get_info(dev)  # cudaMemGetInfo: sees ANY device allocation, the library's own included
    rc = lib.ia_gather_meandiff_static(cptr.data_ptr(), T, 0, 0, C, bptr.data_ptr(), T, 0, 0, Nb, L, H,
                                       kernels.MODES[mode], int(normalize), out32.data_ptr(), outr.data_ptr(),
                                       outn.data_ptr(), kernels._dt(conc[0]), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.ia_error_string(rc)
    assert torch.cuda.memory_allocated(dev) == before  # the call went through no torch allocation
    torch.cuda.synchronize()
    if check_device_free:
        # ... and through no cudaMalloc of its own: the device's free memory is what it was (checked on calls after
        # the first one on this device, whose set-up may make the driver load the kernels' module)
        assert torch.cuda.mem_get_info(dev)[0] == free_before
    return out32, outr, outn


def test_static_scratch_entry_point_matches_workspace_entry_point_twice():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import _lib, kernels

    lib = _lib.load()
    dev = torch.device("cuda:0")
    # call 1: many layers, chunked baseline (> 256 rows) so the column-sum scratch is exercised too
    # call 2: a different, smaller shape with normalisation — stale partials from call 1 would corrupt it
    cases = [((6, 50, 300, 4096, 3, 1), "baseline", False), ((2, 7, 100, 896, 2, 2), "baseline", True),
             ((3, 300, 40, 1024, 2, 3), "posneg", True)]
    for (shape, mode, normalize) in cases:
        conc, base = _case(dev, *shape)
        got = _run_static(lib, conc, base, mode, normalize, kernels)
        want = kernels.gather_meandiff(conc, base, mode, normalize)
        torch.cuda.synchronize()
        for a, b, name in zip(got, want, ("f32", "rounded", "norm")):
            assert not torch.isnan(a.float()).any(), name
            assert torch.equal(a, b), (shape, mode, name)
    # the same request again: a repeat call over the used scratch gives the same bits
    again = _run_static(lib, conc, base, mode, normalize, kernels, check_device_free=True)
    assert all(torch.equal(a, b) for a, b in zip(again, want))


def test_static_scratch_rejects_requests_beyond_its_capacity():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from introspective_awareness_b200 import _lib, kernels

    lib = _lib.load()
    # 128 layers x 5120 columns needs ~5.5 MiB of column-sum scratch alone: over the 4 MiB array
    need = lib.ia_meandiff_workspace_bytes(128, 50, 100, 5120, 0)
    assert need > _lib.IA_STATIC_SCRATCH_BYTES
    p = torch.zeros(64, dtype=torch.int64, device="cuda:0").data_ptr()
    rc = lib.ia_gather_meandiff_static(p, 4, 0, 0, 50, p, 4, 0, 0, 100, 128, 5120, 0, 0, 0, p, 0, 0, 0)
    assert rc == -6 and b"maximum" in lib.ia_error_string(rc)
