#!/usr/bin/env python
"""Collective timings for the view-parallel exchange at the bench sizes (torchrun, one rank per GPU)."""
import os, sys, time
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
P = 3_000_000

def timeit(name, fn, iters=20, nbytes=None):
    for _ in range(5):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        extra = f"  {nbytes / 1e6 / ms.item():8.1f} GB/s (payload/ms)" if nbytes else ""
        print(f"{name:46s} {ms.item():8.3f} ms{extra}", flush=True)

dcol = torch.randn(P, 3, device=dev); all_dcol = torch.empty(world, P, 3, device=dev)
flat = torch.randn(11 * P, device=dev); full = torch.randn(59 * P, device=dev)
timeit(f"all_gather [P,3] x{world} (36 MB/rank)", lambda: dist.all_gather_into_tensor(all_dcol, dcol), nbytes=world * P * 12)
timeit("all_reduce 11P f32 (132 MB)", lambda: dist.all_reduce(flat), nbytes=11 * P * 4)
timeit("all_reduce 59P f32 (708 MB)", lambda: dist.all_reduce(full), nbytes=59 * P * 4)
h = torch.randn(11 * P, device=dev).half()
timeit("all_reduce 11P f16 (66 MB)", lambda: dist.all_reduce(h), nbytes=11 * P * 2)
def both():
    w1 = dist.all_gather_into_tensor(all_dcol, dcol, async_op=True); w2 = dist.all_reduce(flat, async_op=True); w1.wait(); w2.wait()
timeit("all_gather + all_reduce 11P back to back", both)
# symmetric memory availability
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(P * 3, dtype=torch.float32, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD)
    ptrs = list(hdl.buffer_ptrs)
    if rank == 0:
        print("symm_mem ok: world", hdl.world_size, "ptrs", [hex(p) for p in ptrs[:3]], "multicast", hex(getattr(hdl, "multicast_ptr", 0) or 0), flush=True)
    t.copy_(dcol.view(-1))
    hdl.barrier()
    peer = hdl.get_buffer((rank + 1) % world, (P * 3,), torch.float32)
    dst = torch.empty(P * 3, device=dev)
    timeit("peer copy 36 MB (symm_mem get_buffer)", lambda: dst.copy_(peer), nbytes=P * 12)
    def pull_all():
        for r in range(world):
            if r != rank:
                all_dcol[r].view(-1).copy_(hdl.get_buffer(r, (P * 3,), torch.float32))
    timeit(f"pull all peers {world - 1} x 36 MB (copies)", pull_all, nbytes=(world - 1) * P * 12)
    hdl.barrier()
    gname = dist.group.WORLD.group_name
    fs = symm.empty(11 * P, dtype=torch.float32, device=dev); symm.rendezvous(fs, dist.group.WORLD); fs.copy_(flat)
    for name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            timeit(f"symm_mem.{name} 11P f32 (132 MB)", lambda: op(fs, "sum", gname), nbytes=11 * P * 4)
        except Exception as e:
            if rank == 0: print(name, "failed:", repr(e)[:200], flush=True)
    try:
        ag_out = symm.empty(world * P * 3, dtype=torch.float32, device=dev); symm.rendezvous(ag_out, dist.group.WORLD)
        timeit("symm_mem.multimem_all_gather_out 36 MB/rank", lambda: torch.ops.symm_mem.multimem_all_gather_out(t, gname, ag_out), nbytes=world * P * 12)
    except Exception as e:
        if rank == 0: print("multimem_all_gather_out failed:", repr(e)[:200], flush=True)
    # fused all-gather -> SH rebuild: the rebuild kernel reads every peer's [P,3] colour gradients over NVLink
    import importlib
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    pkg = importlib.import_module("3dgs_study_b200"); pkg.ensure_on_path()
    import diff_gaussian_rasterization as dgr
    means = torch.randn(P, 3, device=dev); cp = torch.randn(world, 3, device=dev) * 5
    dsh = torch.empty(P, 16, 3, device=dev)
    timeit("sh_grad_from_views, LOCAL copies of 8 views", lambda: dgr._C.sh_grad_from_views(means, [all_dcol[v] for v in range(world)], cp, 16, 3, out=dsh), nbytes=P * 192)
    timeit("sh_grad_from_views, PEER pointers (fused all-gather)", lambda: dgr._C.sh_grad_from_views(means, ptrs, cp, 16, 3, out=dsh), nbytes=(world - 1) * P * 12)
    def fused_with_barriers():
        hdl.barrier(channel=1); dgr._C.sh_grad_from_views(means, ptrs, cp, 16, 3, out=dsh); hdl.barrier(channel=2)
    timeit("  same + symm_mem barriers before/after", fused_with_barriers)
except Exception as e:
    if rank == 0:
        print("symm_mem unavailable:", repr(e)[:300], flush=True)
dist.barrier(); dist.destroy_process_group()
