"""Numerics + timing of the sharded-data-parallel peer-memory kernels (run under torchrun, >= 2 GPUs):

  gather      smpk_sdp_gather (striped all-gather by peer loads)        vs NCCL all_gather_into_tensor
  gemm+gather smpk_gemm_comm mode 3 (unit all-gather inside the QKV GEMM) vs NCCL all-gather then GEMM
  reduce      smpk_sdp_reduce_adam mode 0 / mode 1                       vs NCCL reduce_scatter (+ multi-tensor Adam)

    torchrun --nproc-per-node 2 tools/sdp_fused_check.py [--hidden 6144] [--tokens 8192]
"""
import argparse
import ctypes
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from smdistributed_modelparallel_b200.ops import _native as N  # noqa: E402
from smdistributed_modelparallel_b200.ops import multi_tensor as MT  # noqa: E402
from smdistributed_modelparallel_b200.ops.gemm import gemm  # noqa: E402
from smdistributed_modelparallel_b200.parallel.fused_tp import MAX_RANKS, SymmetricBuffer, _CommDesc, _ptr_array  # noqa: E402


def timeit(fn, iters=10, warm=3, label="timed loop"):
    """ms per call, max over ranks; the loop runs ~0.25 s (same count on every rank) and rank 0 prints the clock
    record taken during it (SM clock, throttle reasons)."""
    import contextlib

    from tools._clocks import calibrated_iters, clocks
    for _ in range(warm):
        fn()
    n_it = torch.tensor([calibrated_iters(fn, floor=iters)], device="cuda")
    dist.all_reduce(n_it, op=dist.ReduceOp.MIN)
    iters = int(n_it.item())
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with (clocks(label) if dist.get_rank() == 0 else contextlib.nullcontext()):
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hidden", type=int, default=6144)
    ap.add_argument("--tokens", type=int, default=8192)
    ap.add_argument("--small", action="store_true")
    args = ap.parse_args()
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
    rank, n = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dev = torch.device("cuda")
    pg = dist.group.WORLD
    H = 256 if args.small else args.hidden
    T = 512 if args.small else args.tokens
    K, Nq = H, 3 * H
    unit_elems = 12 * H * H                       # one transformer layer: qkv 3H^2, dense H^2, mlp 8H^2
    piece = 8 * K
    mult = n * piece
    numel = (unit_elems + mult - 1) // mult * mult
    per, ppr = numel // n, numel // n // piece
    rows_per_rank = per // K
    if rank == 0:
        print(f"[sdp check] n={n} H={H} T={T} unit {numel / 1e6:.1f} M elems ({numel * 2 / 2**20:.0f} MiB bf16) "
              f"piece {piece * 2 // 1024} KiB, {ppr} pieces/rank")
    torch.manual_seed(7)
    full_ref = (torch.randn(numel, device=dev, dtype=torch.float32) * 0.02).to(torch.bfloat16)   # same on every rank
    shards = SymmetricBuffer(per * 2, None, rank, None, pg=pg)
    flags = SymmetricBuffer(8 * MAX_RANKS * 4, None, rank, None, pg=pg)
    shard = shards.view(torch.bfloat16, per)
    shard.copy_(full_ref.view(ppr, n, piece)[:, rank].reshape(-1))
    blk_flags = torch.zeros(numel // K // 64 + 8, dtype=torch.int32, device=dev)
    stream = N.stream_ptr(dev)
    peers0 = _ptr_array([flags.peer_ptrs[j] for j in range(n)])
    N.call("smpk_flags_signal", stream, ctypes.cast(peers0, ctypes.c_void_p), n, rank, 1)
    torch.cuda.synchronize()
    dist.barrier()
    ok = True

    # ---------------- stand-alone gather
    full = torch.zeros(numel, dtype=torch.bfloat16, device=dev)
    ps = _ptr_array(shards.peer_ptrs)

    def gather():
        N.call("smpk_sdp_gather", stream, ctypes.cast(ps, ctypes.c_void_p), full.data_ptr(), piece * 2, ppr, flags.ptr, 1,
               n, rank, 0)

    gather()
    torch.cuda.synchronize()
    exact = torch.equal(full, full_ref)
    ok &= exact
    contiguous_shard = full_ref[rank * per:(rank + 1) * per].clone()
    full2 = torch.empty_like(full)
    t_g = timeit(gather)
    t_n = timeit(lambda: dist.all_gather_into_tensor(full2, contiguous_shard))
    gb = numel * 2 * (n - 1) / n / 1e9
    if rank == 0:
        print(f"gather exact={exact}: peer-load {t_g:.3f} ms ({gb / t_g * 1e3:.0f} GB/s in) | NCCL all_gather {t_n:.3f} ms "
              f"({gb / t_n * 1e3:.0f} GB/s)")

    # ---------------- gather fused into the QKV GEMM
    x = (torch.randn(T, K, device=dev) * 0.5).to(torch.bfloat16)
    W_ref = full_ref[:Nq * K].view(Nq, K)
    y = torch.empty(T, Nq, dtype=torch.bfloat16, device=dev)

    def fused():
        d = _CommDesc()
        d.mode, d.nranks, d.rank, d.rows_per_rank, d.comm_ctas, d.epoch = 3, n, rank, rows_per_rank, 0, 1
        for j in range(n):
            d.peer_src[j] = shards.peer_ptrs[j]
        d.local_buf = full.data_ptr()
        d.blk_flags = blk_flags.data_ptr()
        d.ready_flags = flags.ptr
        N.call("smpk_gemm_comm", stream, N.dtype_code(x.dtype), N.dtype_code(y.dtype), T, Nq, K, x.data_ptr(), K, 0,
               full.data_ptr(), K, 0, y.data_ptr(), Nq, None, 0, None, 0, 0, 0.0, None, 1.0, 0, 0, ctypes.byref(d))

    full.zero_()
    fused()
    torch.cuda.synchronize()
    ref = x.float() @ W_ref.float().t()
    rel = ((y.float() - ref).norm() / ref.norm()).item()
    exact2 = torch.equal(full, full_ref)
    ok &= exact2 and rel < 1e-2
    Wg = full2[:Nq * K].view(Nq, K)

    def unfused():
        dist.all_gather_into_tensor(full2, contiguous_shard)
        gemm(x, Wg, out=y)

    t_f = timeit(fused)
    t_u = timeit(unfused)
    t_gemm = timeit(lambda: gemm(x, Wg, out=y))
    fl = 2.0 * T * Nq * K
    if rank == 0:
        print(f"gemm+gather rel={rel:.2e} unit exact={exact2}: fused {t_f:.3f} ms | NCCL all_gather + GEMM {t_u:.3f} ms | "
              f"GEMM alone {t_gemm:.3f} ms ({fl / t_gemm / 1e9:.0f} TF) | roofline max(gemm, bytes/link) = "
              f"{max(t_gemm, gb / 900 * 1e3):.3f} ms -> {max(t_gemm, gb / 900 * 1e3) / t_f * 100:.0f} %")

    # ---------------- reduce-scatter (+ Adam)
    grads = SymmetricBuffer(numel * 4, None, rank, None, pg=pg)
    g = grads.view(torch.float32, numel)
    torch.manual_seed(100 + rank)
    g.copy_(torch.randn(numel, device=dev) * 1e-2)
    torch.cuda.synchronize()
    dist.barrier()
    tot = g.clone()
    dist.all_reduce(tot)
    mine = tot.view(ppr, n, piece)[:, rank].reshape(-1)
    grad_shard = torch.zeros(per, dtype=torch.float32, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    pgp = _ptr_array(grads.peer_ptrs)
    pcons = _ptr_array([flags.peer_ptrs[j] + 4 * MAX_RANKS * 2 for j in range(n)])
    pready = _ptr_array([flags.peer_ptrs[j] + 4 * MAX_RANKS * 1 for j in range(n)])
    ep = [0]

    def reduce_only(beta=0.0):
        ep[0] += 1
        N.call("smpk_flags_signal", stream, ctypes.cast(pready, ctypes.c_void_p), n, rank, ep[0])
        N.call("smpk_sdp_reduce_adam", stream, ctypes.cast(pgp, ctypes.c_void_p), grad_shard.data_ptr(), None, None, None,
               None, 0, piece, ppr, flags.ptr + 4 * MAX_RANKS, ctypes.cast(pcons, ctypes.c_void_p), counter.data_ptr(),
               ep[0], n, rank, 1.0, beta, 0, 0.0, 0.0, 0.0, 0.0, 0.0, 1, 0, 0, 1.0, 0)
        # the slot may only be rewritten once every peer has read it
        N.call("smpk_flags_wait", stream, flags.ptr + 4 * MAX_RANKS * 2, n, -1, ep[0])

    reduce_only()
    torch.cuda.synchronize()
    err = (grad_shard - mine).abs().max().item()
    ok &= err < 1e-5
    master = shard.float().clone()
    m1, v1 = torch.zeros_like(master), torch.zeros_like(master)
    rmaster, rm, rv = master.clone(), m1.clone(), v1.clone()
    lowp = torch.empty(per, dtype=torch.bfloat16, device=dev)
    step = [0]

    def reduce_adam():
        ep[0] += 1
        step[0] += 1
        N.call("smpk_flags_signal", stream, ctypes.cast(pready, ctypes.c_void_p), n, rank, ep[0])
        N.call("smpk_sdp_reduce_adam", stream, ctypes.cast(pgp, ctypes.c_void_p), grad_shard.data_ptr(), master.data_ptr(),
               m1.data_ptr(), v1.data_ptr(), lowp.data_ptr(), N.dtype_code(lowp.dtype), piece, ppr, flags.ptr + 4 * MAX_RANKS,
               ctypes.cast(pcons, ctypes.c_void_p), counter.data_ptr(), ep[0], n, rank, 1.0 / n, 0.0, 1, 1e-3, 0.9, 0.95,
               1e-8, 0.1, step[0], 1, 1, 1.0, 0)
        N.call("smpk_flags_wait", stream, flags.ptr + 4 * MAX_RANKS * 2, n, -1, ep[0])

    reduce_adam()
    torch.cuda.synchronize()
    gref = mine / n
    rm.mul_(0.9).add_(gref, alpha=0.1)
    rv.mul_(0.95).addcmul_(gref, gref, value=0.05)
    upd = (rm / (1 - 0.9)) / ((rv / (1 - 0.95)).sqrt() + 1e-8) + 0.1 * rmaster
    rmaster.add_(upd, alpha=-1e-3)
    aerr = (master - rmaster).abs().max().item()
    lerr = (lowp.float() - rmaster).abs().max().item()
    ok &= aerr < 1e-5 and lerr < 1e-2
    tmp = torch.empty(per, dtype=torch.float32, device=dev)
    gcontig = g.clone()
    lists = [[tmp], [master], [m1], [v1], [lowp]]

    def nccl_rs_adam():
        dist.reduce_scatter_tensor(tmp, gcontig)
        MT.multi_tensor_adam(MT.DEFAULT_CHUNK, None, lists, 1e-3, 0.9, 0.95, 1e-8, 5, 1, True, 0.1, 1.0 / n, None)

    t_r = timeit(reduce_only)
    t_ra = timeit(reduce_adam)
    t_nr = timeit(lambda: dist.reduce_scatter_tensor(tmp, gcontig))
    t_na = timeit(nccl_rs_adam)
    rb = per * 4 * (n - 1) / 1e9                          # bytes pulled over NVLink per rank
    hb = per * (4 * n + 4 * 6 + 2) / 1e9                  # HBM traffic of the fused RS+Adam per rank (local slot incl.)
    roof = max(rb / 900, hb / 6500) * 1e3
    if rank == 0:
        print(f"reduce err={err:.1e} adam err={aerr:.1e} lowp err={lerr:.1e}: peer-load RS {t_r:.3f} ms | RS+Adam fused "
              f"{t_ra:.3f} ms | NCCL RS {t_nr:.3f} ms | NCCL RS + Adam {t_na:.3f} ms | roofline {roof:.3f} ms -> "
              f"{roof / t_ra * 100:.0f} %")
        print("NUMERICS PASS" if ok else "NUMERICS FAIL")
    okt = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(okt) == 1 else 1)


if __name__ == "__main__":
    main()
