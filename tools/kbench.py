"""Kernel micro-benchmarks: K2 / K1 achieved HBM GB/s over the SURVEY.md §8(d) shape ladder.

CUDA-event timing on the launching stream, >= 3 warm-ups, L2 flushed between timed launches for
shapes smaller than L2 (a 256 MB write), one JSON line per shape.  Run on the B200 box:
    python tools/kbench.py [--quick] [--impl regs|tma] > gpurun_out/kbench.jsonl
"""

from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from introspective_awareness_b200 import _lib, kernels  # noqa: E402

PEAK_FALLBACK = 6650.0


def peak_gbs():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    except Exception:
        return PEAK_FALLBACK, "fallback"


def time_launch(fn, reps, flush):
    """Average device time (ms) of fn() over reps launches, each preceded by an L2 flush when asked."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    for a, b in ev:
        total += a.elapsed_time(b)
    return total / reps


def bench_k2(rows, T, H, variant, impl, reps, flush, dtype=torch.bfloat16, add=False, add_staged=False):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(rows, T, H, device=dev, generator=g).to(dtype)
    y = torch.empty_like(h)
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(dtype)
    w = torch.ones(H, device=dev, dtype=dtype)
    row_vec = (torch.arange(rows, device=dev) % 50).to(torch.int32)
    alpha = torch.full((rows,), 4.0, device=dev)
    t0 = torch.zeros(rows, device=dev, dtype=torch.int32)
    e = h.element_size()

    lib = _lib.load()
    flags = _lib.IA_STEER_ALL | {"auto": 0, "tma": _lib.IA_IMPL_TMA, "regs": _lib.IA_IMPL_REGS, "stream": _lib.IA_IMPL_STREAM}[impl]
    if add_staged:
        flags |= _lib.IA_IMPL_ADD_STAGED
    args = (h.data_ptr(), y.data_ptr(), vt.data_ptr(), row_vec.data_ptr(), alpha.data_ptr(), t0.data_ptr(),
            w.data_ptr(), 1e-5, rows, T, H, 0, kernels._dt(h), kernels._dt(vt), kernels.VARIANTS[variant], flags,
            torch.cuda.current_stream().cuda_stream)

    branch = (torch.randn(rows, T, H, device=dev, generator=g) * 1e-3).to(dtype) if add else None
    if add:
        # residual-add form: h <- h + branch first; five streams per vector (h, branch, table row, h', y)
        args = args[:1] + (branch.data_ptr(),) + args[1:]

    def fn():  # the raw C-ABI call: no per-call Python tensor work inside the timed region
        rc = lib.ia_add_inject_rmsnorm(*args) if add else lib.ia_inject_rmsnorm(*args)
        assert rc == 0, rc

    ms = time_launch(fn, reps, flush)
    algo = rows * T * H * e * (5 if add else 4) + H * e
    dram = rows * T * H * e * (4 if add else 3)  # the table row is an L2 hit: bytes DRAM must move
    return dict(kernel="k2_add_inject_rmsnorm" if add else "k2_inject_rmsnorm", impl=impl, variant=variant, rows=rows, T=T,
                H=H, bytes=algo, dram_bytes=dram, us=ms * 1e3, gbs=algo / ms / 1e6, dram_gbs=dram / ms / 1e6)


def bench_tail(rows, H, impl, reps, flush, dtype=torch.bfloat16):
    """Gemma-2 block tail (ia_gemma2_tail_inject_rmsnorm): five streams per vector."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    res = torch.randn(rows, 1, H, device=dev, generator=g).to(dtype)
    mlp = torch.randn(rows, 1, H, device=dev, generator=g).to(dtype)
    y = torch.empty_like(res)
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(dtype)
    w = torch.zeros(H, device=dev, dtype=dtype)
    row_vec = (torch.arange(rows, device=dev) % 50).to(torch.int32)
    alpha = torch.full((rows,), 4.0, device=dev)
    t0 = torch.zeros(rows, device=dev, dtype=torch.int32)
    lib = _lib.load()
    flags = _lib.IA_STEER_ALL | {"auto": 0, "tma": _lib.IA_IMPL_TMA, "regs": _lib.IA_IMPL_REGS, "stream": _lib.IA_IMPL_STREAM}[impl]
    args = (mlp.data_ptr(), res.data_ptr(), y.data_ptr(), vt.data_ptr(), row_vec.data_ptr(), alpha.data_ptr(), t0.data_ptr(),
            w.data_ptr(), w.data_ptr(), 1e-6, rows, 1, H, 0, kernels._dt(res), kernels._dt(vt), flags,
            torch.cuda.current_stream().cuda_stream)

    def fn():
        rc = lib.ia_gemma2_tail_inject_rmsnorm(*args)
        assert rc == 0, rc

    ms = time_launch(fn, reps, flush)
    algo = rows * H * res.element_size() * 5 + 2 * H * res.element_size()
    dram = rows * H * res.element_size() * 4
    return dict(kernel="k2_gemma2_tail", impl=impl, rows=rows, T=1, H=H, bytes=algo, dram_bytes=dram, us=ms * 1e3,
                gbs=algo / ms / 1e6, dram_gbs=dram / ms / 1e6)


def bench_k2_graph(rows, H, form, reps, n_in_graph=64, n_sets=8, dtype=torch.bfloat16, impl="auto"):
    """Small shapes as the decode loop runs them: `n_in_graph` launches captured back to back in ONE CUDA graph (rotating
    over `n_sets` buffer sets, as consecutive layers do), replayed; us = replay time / launches.  The cold / hot figures of
    the ladder at these sizes measure the launch, not the kernel ("the measured us inside a CUDA graph on
    the small ones")."""
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    lib = _lib.load()
    vt = (torch.randn(50, H, device=dev, generator=g) * 0.1).to(dtype)
    w = torch.ones(H, device=dev, dtype=dtype) if form != "tail" else torch.zeros(H, device=dev, dtype=dtype)
    row_vec = (torch.arange(rows, device=dev) % 50).to(torch.int32)
    alpha = torch.full((rows,), 4.0, device=dev)
    t0 = torch.zeros(rows, device=dev, dtype=torch.int32)
    sets = [[torch.randn(rows, 1, H, device=dev, generator=g).to(dtype) for _ in range(3)] for _ in range(n_sets)]
    flags = _lib.IA_STEER_ALL | {"auto": 0, "tma": _lib.IA_IMPL_TMA, "regs": _lib.IA_IMPL_REGS, "stream": _lib.IA_IMPL_STREAM}[impl]

    def launch(k, stream):
        h, b, y = sets[k % n_sets]
        if form == "plain":
            rc = lib.ia_inject_rmsnorm(h.data_ptr(), y.data_ptr(), vt.data_ptr(), row_vec.data_ptr(), alpha.data_ptr(), t0.data_ptr(),
                                       w.data_ptr(), 1e-5, rows, 1, H, 0, kernels._dt(h), kernels._dt(vt), kernels.VARIANTS["llama"], flags, stream)
        elif form == "add":
            rc = lib.ia_add_inject_rmsnorm(h.data_ptr(), b.data_ptr(), y.data_ptr(), vt.data_ptr(), row_vec.data_ptr(), alpha.data_ptr(),
                                           t0.data_ptr(), w.data_ptr(), 1e-5, rows, 1, H, 0, kernels._dt(h), kernels._dt(vt),
                                           kernels.VARIANTS["llama"], flags, stream)
        else:
            rc = lib.ia_gemma2_tail_inject_rmsnorm(b.data_ptr(), h.data_ptr(), y.data_ptr(), vt.data_ptr(), row_vec.data_ptr(), alpha.data_ptr(),
                                                   t0.data_ptr(), w.data_ptr(), w.data_ptr(), 1e-6, rows, 1, H, 0, kernels._dt(h), kernels._dt(vt),
                                                   flags, stream)
        assert rc == 0, rc

    launch(0, torch.cuda.current_stream().cuda_stream)  # set-up calls (shared-memory opt-in) happen outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        st = torch.cuda.current_stream().cuda_stream
        for k in range(n_in_graph):
            launch(k, st)
    ms = time_launch(graph.replay, reps, None)
    e = sets[0][0].element_size()
    streams = {"plain": 4, "add": 5, "tail": 5}[form]
    algo = rows * H * e * streams + H * e
    us = ms * 1e3 / n_in_graph
    return dict(kernel={"plain": "k2_inject_rmsnorm", "add": "k2_add_inject_rmsnorm", "tail": "k2_gemma2_tail"}[form], impl=impl,
                rows=rows, T=1, H=H, bytes=algo, us=us, gbs=algo / us / 1e3, cache="graph", launches_per_replay=n_in_graph,
                note="back-to-back launches inside one CUDA graph replay, L2-resident operands (the decode loop's situation)")


def bench_copy(nbytes, reps, flush):
    """torch's device copy moving `nbytes` in total (half read, half written), timed the same way: what the copy
    that defines MEASURED_PEAKS.json's hbm_gbs achieves at THIS size (launch ramp and tail included)."""
    n = nbytes // 4  # bf16 elements each way
    a = torch.empty(n, dtype=torch.bfloat16, device="cuda:0").normal_()
    b = torch.empty_like(a)
    ms = time_launch(lambda: b.copy_(a), reps, flush)
    return dict(kernel="torch_copy", bytes=2 * n * 2, us=ms * 1e3, gbs=2 * n * 2 / ms / 1e6)


def bench_k1_compact(n_layers, C, Nb, H, reps, flush, dtype=torch.bfloat16, want_norm=False):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    ca = torch.randn(n_layers, C, H, device=dev, generator=g).to(dtype)
    ba = torch.randn(n_layers, Nb, H, device=dev, generator=g).to(dtype)
    e = ca.element_size()
    ws = torch.empty(kernels.workspace_bytes(n_layers, C, Nb, H) + 256, dtype=torch.uint8, device=dev)
    out_r = torch.empty(n_layers, C, H, dtype=dtype, device=dev)
    out_n = torch.empty(n_layers, C, dtype=torch.float32, device=dev)
    lib = _lib.load()
    args = (ca.data_ptr(), C, ba.data_ptr(), Nb, n_layers, H, _lib.IA_MODE_BASELINE, 0, 0, out_r.data_ptr(),
            0 if not want_norm else out_n.data_ptr(), ws.data_ptr(), ws.numel(), kernels._dt(ca), torch.cuda.current_stream().cuda_stream)

    def fn():
        rc = lib.ia_meandiff_compact(*args)
        assert rc == 0, rc

    ms = time_launch(fn, reps, flush)
    algo = n_layers * ((C + Nb) * H * e + C * H * e)
    return dict(kernel="k1_meandiff_compact", n_layers=n_layers, C=C, Nb=Nb,
                H=H, norm_pass=want_norm, bytes=algo, us=ms * 1e3, gbs=algo / ms / 1e6)


def bench_k1(n_layers, C, Nb, H, reps, flush, dtype=torch.bfloat16, want_norm=False):
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    T = 1
    cl = [torch.randn(C, T, H, device=dev, generator=g).to(dtype) for _ in range(n_layers)]
    bl = [torch.randn(Nb, T, H, device=dev, generator=g).to(dtype) for _ in range(n_layers)]
    e = cl[0].element_size()
    ws = torch.empty(kernels.workspace_bytes(n_layers, C, Nb, H) + 256, dtype=torch.uint8, device=dev)
    cptr, bptr = kernels._ptr_table(cl, "c"), kernels._ptr_table(bl, "b")
    out_r = torch.empty(n_layers, C, H, dtype=dtype, device=dev)
    out_n = torch.empty(n_layers, C, dtype=torch.float32, device=dev)
    lib = _lib.load()
    # out_norm = NULL: the production call (the driver never normalises) needs no norm pass
    args = (cptr.data_ptr(), T, 0, 0, C, bptr.data_ptr(), T, 0, 0, Nb, n_layers, H, _lib.IA_MODE_BASELINE, 0, 0,
            out_r.data_ptr(), 0 if not want_norm else out_n.data_ptr(), ws.data_ptr(), ws.numel(), kernels._dt(cl[0]),
            torch.cuda.current_stream().cuda_stream)

    def fn():
        rc = lib.ia_gather_meandiff(*args)
        assert rc == 0, rc

    ms = time_launch(fn, reps, flush)
    algo = n_layers * ((C + Nb) * H * e + C * H * e)
    return dict(kernel="k1_gather_meandiff", n_layers=n_layers, C=C, Nb=Nb, H=H, norm_pass=want_norm, bytes=algo, us=ms * 1e3,
                gbs=algo / ms / 1e6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--impl", default="auto", choices=["auto", "regs", "tma", "stream"])
    ap.add_argument("--r2", action="store_true", help="round-2 ladder: streaming vs staged K2 (plain / add / tail), torch copy at "
                                                      "the same sizes, compact K1")
    ap.add_argument("--only-form", default="plain", choices=["plain", "add", "tail"], help="with --only-rows")
    ap.add_argument("--only-h", type=int, default=4096, help="with --only-rows")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--only-rows", type=int, default=0, help="run a single cold K2 shape (for ncu --set full)")
    ap.add_argument("--only-k1", action="store_true", help="run only the K1 ladder (and, with --r2, the graph-timed small K2 shapes)")
    ap.add_argument("--add", action="store_true", help="K2 in its residual-add form (ia_add_inject_rmsnorm)")
    ap.add_argument("--add-staged", action="store_true", help="with --add: force the bulk-copied branch row (IA_IMPL_ADD_STAGED)")
    args = ap.parse_args()
    peak, src = peak_gbs()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda:0")
    if args.r2:
        def emit(r, **extra):
            r.update(cache="cold", frac=r["gbs"] / peak, peak=peak, peak_src=src, **extra)
            if "dram_gbs" in r:
                r["dram_frac"] = r["dram_gbs"] / peak
            print(json.dumps(r), flush=True)

        for rows in (() if args.only_k1 else (3000, 8192, 25600, 262144)):
            for H in (4096, 5120, 3584):
                for impl in ("stream", "tma"):
                    emit(bench_k2(rows, 1, H, "gemma" if H == 3584 else "llama", impl, args.reps, flush))
                    if H != 3584:
                        emit(bench_k2(rows, 1, H, "llama", impl, args.reps, flush, add=True))
                    else:
                        emit(bench_tail(rows, H, impl, args.reps, flush))
            for nbytes in (rows * 4096 * 2 * 3, rows * 4096 * 2 * 4):
                emit(bench_copy(nbytes, args.reps, flush))
        for rows in (200, 3000):
            for H, form in ((4096, "plain"), (4096, "add"), (5120, "add"), (3584, "tail")):
                r = bench_k2_graph(rows, H, form, args.reps)
                r.update(frac=r["gbs"] / peak, peak=peak, peak_src=src)
                print(json.dumps(r), flush=True)
        for (L, C, Nb) in [(1, 50, 100), (32, 50, 100), (64, 50, 100), (1, 50, 16334), (1, 50, 262094)]:
            for want_norm in (False, True):
                emit(bench_k1_compact(L, C, Nb, 4096, args.reps, flush, want_norm=want_norm))
            emit(bench_k1(L, C, Nb, 4096, args.reps, flush))
        return
    if args.only_rows and args.only_form == "tail":
        r = bench_tail(args.only_rows, args.only_h, args.impl, args.reps, flush)
        r.update(cache="cold", frac=r["gbs"] / peak, peak=peak, peak_src=src)
        print(json.dumps(r), flush=True)
        return
    if args.only_rows:
        args.add = args.add or args.only_form == "add"
        r = bench_k2(args.only_rows, 1, args.only_h, "llama", args.impl, args.reps, flush, add=args.add, add_staged=args.add_staged)
        r["add_staged"] = args.add_staged
        r.update(cache="cold", frac=r["gbs"] / peak, peak=peak, peak_src=src)
        print(json.dumps(r), flush=True)
        return
    rows_ladder = [200, 3000, 25600] if args.quick else [200, 3000, 25600, 262144]
    hs = [4096] if args.quick else [896, 3584, 4096, 5120]
    if args.only_k1:
        hs = []
    for H in hs:
        for rows in rows_ladder:
            if rows * H * 2 * 2 > 40e9:
                continue
            for cache, fl in (("cold", flush), ("hot", None)):
                if cache == "hot" and rows * H * 2 * 3 > 100e6:
                    continue  # larger than L2 anyway
                for with_add in ((False, True) if H != 3584 else (False,)):  # Gemma-2 keeps the HF block
                    r = bench_k2(rows, 1, H, "gemma" if H == 3584 else "llama", args.impl, args.reps, fl, add=with_add)
                    r.update(cache=cache, frac=r["gbs"] / peak, peak=peak, peak_src=src)
                    print(json.dumps(r), flush=True)
    if not args.only_k1:
        for rows in ([25600] if args.quick else [200, 25600, 262144]):
            for impl in ("auto", "regs"):
                r = bench_tail(rows, 3584, impl, args.reps, flush)
                r.update(cache="cold", frac=r["gbs"] / peak, peak=peak, peak_src=src)
                print(json.dumps(r), flush=True)
    for (L, C, Nb) in ([(32, 50, 100)] if args.quick else [(1, 50, 100), (32, 50, 100), (1, 50, 16334), (1, 50, 262094)]):
        r = bench_k1(L, C, Nb, 4096, args.reps, flush)
        r.update(frac=r["gbs"] / peak, peak=peak, peak_src=src)
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
