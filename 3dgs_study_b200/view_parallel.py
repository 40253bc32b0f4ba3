"""View-parallel training step (BASELINE.json configs[3]).

The reference renders ONE view per optimizer step on ONE GPU (/root/reference/train.py:86-88,98) and has
no distributed code.  The path shards naturally over camera views: Gaussians replicated, each rank
rasterizes its own views, and the per-Gaussian parameter gradients (59 f32 per Gaussian at SH3:
xyz 3 + SH 48 + opacity 1 + scale 3 + rotation 4, the optimizer groups of
/root/reference/scene/gaussian_model.py:185-192) are sum-allreduced once per step.

Densification side statistics keep the reference's semantics
(/root/reference/scene/gaussian_model.py:576-581, /root/reference/train.py:127): the screen-space
gradient NORM is taken per view and then summed (so it is accumulated locally per view, then
sum-reduced), `denom` counts visible views, `max_radii2D` is max-reduced.

Three exchange modes:
  "packed"      (default for world_size > 1 on CUDA) the per-Gaussian backward writes ONE 64-byte row per Gaussian
                the view touched (geometry gradients + masked colour gradient, b200splat_backward_packed) into a
                symmetric (peer-mapped) buffer; after one barrier every rank READS its peers' rows over NVLink in
                the kernel that adds them into its dense gradients and rebuilds dL/dsh from the colour gradient and
                the camera centre (b200splat_accumulate_packed).  No collective on the data path; the bytes on NVLink
                follow the Gaussians a view actually touched, not P (measured scaling: BASELINE.md section 5).
  "allreduce"   every rank all-reduces the full 59 f32 / Gaussian (708 MB at 3 M SH3).  The plain baseline.
  "factorized"  (the NCCL flavour; default with an injected backend) the SH gradient of a view is the outer product
                Y(dir(mean, campos_v)) (x) dcol_v, so ranks all-gather the [P,3] masked colour gradients
                (12 B per Gaussian and view) plus the camera centres, all-reduce only the remaining
                11 f32 / Gaussian (rotation 4, mean 3, scale 3, opacity 1) as ONE flat buffer, and every rank
                rebuilds the summed dL/dsh locally with b200splat_sh_grad_from_views while the all-reduce is
                still in flight.  Same result (up to fp32 summation order), 3.4x fewer bytes on NVLink at 8 GPUs.

In EVERY mode the densification statistics are rank-local accumulators that `step()` adds to and never clears;
`reduce_densification_stats()` returns their sum / max over the ranks when the caller is about to consume them
(the reference does so every 100 iterations, /root/reference/train.py:130-133) and
`reset_densification_stats()` clears them afterwards, as the reference's densification_postfix does.

Works with any torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests -- the reduction
logic is backend-agnostic and is tested with world_size 2 on CPU with a stand-in rasterize function).
"""
from __future__ import annotations

from typing import Callable, List, NamedTuple, Optional, Sequence

import torch


class GaussianParams(NamedTuple):
    means3D: torch.Tensor    # [P,3]
    shs: torch.Tensor        # [P,M,3]
    opacities: torch.Tensor  # [P,1]
    scales: torch.Tensor     # [P,3]
    rotations: torch.Tensor  # [P,4]


GRAD_FLOATS_PER_GAUSSIAN = lambda M: 3 + 3 * M + 1 + 3 + 4   # noqa: E731  (59 at SH3)


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Rank r renders views {r, r + G, r + 2G, ...}."""
    return list(range(rank, n_views, world_size))


class _CudaBackend:
    """Factorized-mode backend: the CUDA library through `_C` (no autograd: the step owns the gradient buffers)."""

    def __init__(self, step: "ViewParallelStep"):
        import diff_gaussian_rasterization as dgr
        self.C, self.st = dgr._C, step

    def view_fwd_bwd(self, cam, dL, lv, out, on_forward=None, on_colors=None):
        C, st = self.C, self.st
        R, color, radii, geom, binning, img, _ = C.rasterize_gaussians(
            st.bg, lv.means3D, None, lv.opacities, lv.scales, lv.rotations, st.scale_modifier, None,
            cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, cam.image_height,
            cam.image_width, lv.shs, st.sh_degree, cam.camera_center, False, False)
        if on_forward is not None:
            on_forward(color)
        g = C.rasterize_gaussians_backward(
            st.bg, lv.means3D, radii, None, lv.scales, lv.rotations, st.scale_modifier, None, cam.world_view_transform,
            cam.full_proj_transform, cam.tanfovx, cam.tanfovy, dL, lv.shs, st.sh_degree, cam.camera_center, geom, R,
            binning, img, False, opacities=lv.opacities, image_height=cam.image_height, image_width=cam.image_width,
            out={k: v for k, v in out.items() if k != "dL_dcolors"}, early_colors=out["dL_dcolors"],
            after_early_colors=on_colors)
        return color, radii, g[0]

    def sh_grad_from_views(self, means3D, dcols, campos, M, degree):
        # the C ABI takes at most B200SPLAT_MAX_VIEWS (16) views per launch: more (8 ranks x 3 views) go in groups
        out = None
        for lo in range(0, len(dcols), 16):
            part = self.C.sh_grad_from_views(means3D, dcols[lo:lo + 16], campos[lo:lo + 16].contiguous(), M, degree)
            out = part if out is None else out.add_(part)
        return out

    def view_fwd_bwd_packed(self, cam, dL, lv, rows_ptr, count_ptr, dmeans2D, on_forward=None, clean_scratch=None,
                            status_stream=None, stats=None):
        """Forward + backward of one view; the gradients leave as packed rows (sparse exchange).
        stats = (xyz_gradient_accum, denom, max_radii2D) with clean_scratch and status_stream: the densification
        statistics of the view are updated on the way -- the radii part on `status_stream` beside the render backward,
        the gradient norms inside the per-Gaussian backward -- and no dense dL/dmeans2D is produced."""
        C, st = self.C, self.st
        R, color, radii, geom, binning, img, _ = C.rasterize_gaussians(
            st.bg, lv.means3D, None, lv.opacities, lv.scales, lv.rotations, st.scale_modifier, None,
            cam.world_view_transform, cam.full_proj_transform, cam.tanfovx, cam.tanfovy, cam.image_height,
            cam.image_width, lv.shs, st.sh_degree, cam.camera_center, False, False, status_stream=status_stream)
        if on_forward is not None:
            on_forward(color)
        accum = None
        if stats is not None:
            accum, denom, max_radii = stats
            fwd_done = torch.cuda.Event(); fwd_done.record(torch.cuda.current_stream(radii.device))
            status_stream.wait_event(fwd_done)
            with torch.cuda.stream(status_stream):
                C.densify_stats_visible(radii, denom.view(-1), max_radii)
            accum = accum.view(-1)
        C.rasterize_gaussians_backward_packed(
            st.bg, lv.means3D, radii, lv.scales, lv.rotations, st.scale_modifier, cam.world_view_transform,
            cam.full_proj_transform, cam.tanfovx, cam.tanfovy, dL, lv.shs, st.sh_degree, cam.camera_center, geom, R,
            binning, img, lv.opacities, rows_ptr, count_ptr, cam.image_height, cam.image_width, dL_dmeans2D=dmeans2D,
            clean_scratch=clean_scratch, xyz_gradient_accum=accum)
        return color, radii, img


class ViewParallelStep:
    """One data-parallel step: render this rank's views fwd+bwd, accumulate, reduce across ranks."""

    def __init__(self, params: GaussianParams, sh_degree: int, bg: torch.Tensor, world_size: int = 1,
                 process_group=None, rasterize: Optional[Callable] = None, scale_modifier: float = 1.0,
                 exchange: Optional[str] = None, backend=None, peer_memory: Optional[bool] = None):
        self.world_size = world_size
        self.group = process_group
        self.sh_degree = sh_degree
        self.bg = bg
        self.scale_modifier = scale_modifier
        self.leaves = GaussianParams(*[t.detach().clone().requires_grad_(True) for t in params])
        P = self.leaves.means3D.shape[0]
        dev = self.leaves.means3D.device
        self.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
        self.denom = torch.zeros(P, 1, device=dev)
        self.max_radii2D = torch.zeros(P, device=dev)
        self._rasterize = rasterize          # injectable for CPU tests; default = the CUDA operator
        # `backend` (factorized mode): object with view_fwd_bwd(cam, dL, out) and sh_grad_from_views(...);
        # default = the CUDA library through _C, CPU tests inject a torch stand-in
        self._backend = backend
        if exchange is None:
            # N > 1 on CUDA: "auto" (packed rows over peer memory unless the views touch too many Gaussians for them to
            # pay, see _auto_decide); with an injected backend (CPU tests): the factorized exchange over the process
            # group; with an injected rasterizer: plain all-reduce
            if world_size > 1 and rasterize is None:
                exchange = "auto" if (backend is None and dev.type == "cuda") else "factorized"
            else:
                exchange = "allreduce"
        if exchange not in ("allreduce", "factorized", "packed", "auto"):
            raise ValueError("exchange must be 'allreduce', 'factorized', 'packed' or 'auto'")
        # auto: every AUTO_EVERY steps one step runs packed (its row counts are the measurement), the next step takes
        # the largest touched fraction over all ranks (one 4-byte max-all-reduce + host read per AUTO_EVERY steps) and
        # keeps "packed" below AUTO_THRESHOLD, "factorized" above: packed moves 64 B per touched Gaussian and view and
        # pays 15 L2 reductions for each, factorized moves 12 B + 44 B per Gaussian whatever the view touched
        # (measured crossover on 8 x B200: BASELINE.md section 5).
        self._auto = exchange == "auto"
        self.AUTO_EVERY, self.AUTO_THRESHOLD = 64, 0.25
        self._step_no, self._probe_pending = 0, False
        self.exchange = "packed" if self._auto else exchange
        self.last_radii = None
        self._side = None                    # side streams of the factorized / packed exchange
        self._packed = None                  # symmetric buffers of the packed exchange
        self._local_rows = None              # packed rows of a single-rank step (same layout, plain device memory)
        self._dense = None                   # dense gradients of the packed exchange (kept and cleared sparsely from step to step)
        self._clean_scratch = None           # scratch of the packed backward, clean between calls (b200splat_backward_packed_clean)
        # peer_memory (factorized, CUDA, world_size > 1): exchange through NVLink peer / multicast mappings
        # (torch symmetric memory) instead of NCCL -- the SH rebuild kernel then READS every peer's colour gradients
        # itself (all-gather fused into the kernel) and the 11 f32 / Gaussian go through an in-switch (NVLS
        # multimem) all-reduce.  True = require it, None = try and fall back to NCCL, False (default) = NCCL.
        # Measured on 8 x B200 (tools/comm_bench.py, profiles/): the exchange is bound by NVLink wire time either
        # way (all-gather 0.43 ms + all-reduce 0.40 ms vs fused peer rebuild 0.58 ms + multimem all-reduce 0.34 ms)
        # and the NCCL flavour overlaps better with the per-Gaussian backward, so NCCL is the default.
        self.peer_memory = False if peer_memory is None else peer_memory
        self._symm = None

    def _accumulate_stats(self, dmeans2D, radii):
        """Densification statistics of one view (reference semantics: norm per view, then summed)."""
        if dmeans2D.is_cuda and self._rasterize is None and self._backend is None:
            import diff_gaussian_rasterization as dgr
            dgr._C.densify_stats(dmeans2D.contiguous(), radii, self.xyz_gradient_accum.view(-1), self.denom.view(-1),
                                 self.max_radii2D)
            return
        vis = radii > 0
        self.xyz_gradient_accum += torch.norm(dmeans2D[:, :2], dim=-1, keepdim=True) * vis[:, None]
        self.denom += vis[:, None].to(self.denom.dtype)
        self.max_radii2D = torch.maximum(self.max_radii2D, radii.to(self.max_radii2D.dtype))

    # -- rasterizer call ------------------------------------------------------------------------------
    def _render(self, cam, means2D):
        if self._rasterize is not None:
            return self._rasterize(cam, self.leaves, means2D, self.sh_degree, self.bg, self.scale_modifier)
        import diff_gaussian_rasterization as dgr
        rs = dgr.GaussianRasterizationSettings(cam.image_height, cam.image_width, cam.tanfovx, cam.tanfovy, self.bg,
                                               self.scale_modifier, cam.world_view_transform, cam.full_proj_transform,
                                               self.sh_degree, cam.camera_center, False, False)
        lv = self.leaves
        # the training step consumes (color, radii) only, like the reference's call site
        # (/root/reference/gaussian_renderer/__init__.py:98): the inverse-depth channel is not requested
        color, radii, _ = dgr.rasterize_gaussians(lv.means3D, means2D, lv.shs, None, lv.opacities, lv.scales,
                                                  lv.rotations, None, rs, want_invdepth=False)
        return color, radii

    # -- one step -------------------------------------------------------------------------------------
    def step(self, cams: Sequence, dL_images: Sequence[torch.Tensor], reduce: bool = True, on_forward=None) -> torch.Tensor:
        """`on_forward(image)` (optional) is called right after each view's forward has been enqueued, before its
        backward: the hook a caller uses to start consuming the image (e.g. a device->host copy on another stream)."""
        if self._auto and self.world_size > 1 and reduce:
            self._auto_decide(len(cams))
        self._step_no += 1
        if self.exchange == "factorized":
            return self._step_factorized(cams, dL_images, reduce, on_forward)
        if self.exchange == "packed":
            if self.world_size > 1 and reduce and self._packed is None:
                try:
                    self._setup_packed(self.leaves.means3D.shape[0], len(cams), self.leaves.means3D.device)
                except Exception as e:      # no peer mappings on this system: the NCCL flavour of the exchange
                    import warnings
                    warnings.warn(f"packed exchange unavailable ({e!r}); using the factorized NCCL exchange")
                    self.exchange = "factorized"
                    return self._step_factorized(cams, dL_images, reduce, on_forward)
            return self._step_packed(cams, dL_images, reduce, on_forward)
        for t in self.leaves:
            t.grad = None
        color = None
        for cam, dL in zip(cams, dL_images):
            means2D = torch.zeros_like(self.leaves.means3D, requires_grad=True)
            color, radii = self._render(cam, means2D)
            if on_forward is not None:
                on_forward(color.detach())
            torch.autograd.backward([color], [dL])
            with torch.no_grad():
                self._accumulate_stats(means2D.grad, radii)
            self.last_radii = radii
        if reduce and self.world_size > 1:
            self.all_reduce_grads()
        return color.detach()

    def _auto_decide(self, nv):
        """exchange="auto": see __init__.  Collective (every rank calls step() the same number of times)."""
        import torch.distributed as dist
        if self._probe_pending and self._packed is not None:
            pk = self._packed
            P = self.leaves.means3D.shape[0]
            counts = pk["buf"].view(pk["nv"], -1)[:, 0].contiguous().view(torch.int32).max().to(torch.float32).reshape(1)
            dist.all_reduce(counts, op=dist.ReduceOp.MAX, group=self.group)
            self.touched_fraction = float(counts.item()) / max(P, 1)
            self.exchange = "factorized" if self.touched_fraction > self.AUTO_THRESHOLD else "packed"
            self._probe_pending = False
        elif self._step_no % self.AUTO_EVERY == 0:
            self.exchange, self._probe_pending = "packed", True      # this step measures

    # -- packed (sparse, peer-memory) exchange ----------------------------------------------------------------
    # A view touches a few per cent of the Gaussians, and every gradient of an untouched one is exactly zero.  Each
    # rank therefore writes ONE 64-byte row per touched Gaussian (colour gradient x clamp mask, d rotation, d mean,
    # d scale, d opacity) into memory its peers have mapped (torch symmetric memory), and after a barrier every rank
    # runs one kernel that reads all ranks' rows over NVLink and adds them into its dense gradients, rebuilding
    # dL/dsh = Y(dir) (x) dcol on the spot (b200splat_accumulate_packed).  No collective moves gradient data: the
    # exchange IS that kernel.  Bench scene, 8 GPUs: 8 x 14 MB of rows instead of 288 MB all-gather + 132 MB all-reduce.
    def _setup_packed(self, P, nv, dev):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        if self._packed is not None and self._packed["nv"] == nv:
            return self._packed
        group = self.group if self.group is not None else dist.group.WORLD
        stride = 16 * P + 16                                   # words per view: header (count at word 0) + P rows
        buf = symm_mem.empty(nv * stride, dtype=torch.float32, device=dev)
        hd = symm_mem.rendezvous(buf, group)
        buf.zero_()
        self._packed = {"nv": nv, "buf": buf, "hd": hd, "stride": stride, "ptrs": [int(p) for p in hd.buffer_ptrs]}
        return self._packed

    def _step_packed(self, cams, dL_images, reduce, on_forward):
        import torch.distributed as dist
        be = _CudaBackend(self)
        C = be.C
        lv = GaussianParams(*[t.detach() for t in self.leaves])
        P, M = lv.means3D.shape[0], lv.shs.shape[1]
        dev = lv.means3D.device
        G = self.world_size if reduce else 1
        nv = len(cams)
        main = torch.cuda.current_stream(dev)
        if self._side is None:
            self._side = (torch.cuda.Stream(device=dev, priority=-1), torch.cuda.Stream(device=dev))
        comm, side = self._side
        # Dense gradients (the rows are ADDED into them at the end of the step).  They belong to the stepper and live
        # from step to step: zero everywhere except the rows the last step's accumulate calls wrote, and exactly those
        # rows are cleared again (b200splat_clear_accumulated, by the marks the accumulate calls leave) instead of all 0.85 GB
        # -- a full clear next to the render backward costs that kernel 0.1 ms (measured, whichever way it is launched).
        # Nobody else may write into `.grad` of the leaves between steps (reading, scaling in place are fine).
        dn = self._dense
        if dn is None or dn[0].shape[0] != P or dn[1].shape[1] != M or dn[0].device != dev:
            dn = self._dense = (torch.zeros(P, 12, dtype=torch.float32, device=dev),    # (d rotation 4, d mean 3, d scale 3,
                                torch.zeros(P, M, 3, dtype=torch.float32, device=dev),  #  d opacity, 0)
                                torch.zeros(P, dtype=torch.uint8, device=dev))          # marks: Gaussians with a row added
        geom, dsh, marks = dn
        box = {}
        # the sparse clear runs beside the forward's first kernels, on the side stream
        step_start = torch.cuda.Event(); step_start.record(main)
        side.wait_event(step_start)
        with torch.cuda.stream(side):
            C.clear_accumulated(dsh, geom, marks)
            box["cleared"] = torch.cuda.Event(); box["cleared"].record(side)
        stride = 16 * P + 16                                           # words per view: 64-byte header + P rows
        rows_free = None
        if G > 1:
            pk = self._setup_packed(P, nv, dev)
            my = pk["ptrs"][dist.get_rank(self.group)]
            buf = pk["buf"]
            # every rank has read the rows of the previous step (it is past its accumulate calls): they may be overwritten.
            # On the communication stream, beside the forward; the first backward of this step waits for it.
            comm.wait_event(step_start)
            with torch.cuda.stream(comm):
                pk["hd"].barrier(channel=1)
                rows_free = torch.cuda.Event(); rows_free.record(comm)
        else:
            if self._local_rows is None or self._local_rows.numel() != nv * stride:
                self._local_rows = torch.zeros(nv * stride, dtype=torch.float32, device=dev)
            buf = self._local_rows
            my = buf.data_ptr()
        base = [my + 4 * k * stride for k in range(nv)]
        # header of every view: word 0 = row count, words 4..6 = its camera centre, both written by the backward: that is
        # how the peers learn them (b200splat_accumulate_packed with campos = NULL), no collective anywhere in the step
        color = None
        held = []
        for k, (cam, dL) in enumerate(zip(cams, dL_images)):
            def after_forward(img, first=(k == 0)):
                if on_forward is not None:
                    on_forward(img)
                if first and rows_free is not None:
                    main.wait_event(rows_free)           # the first backward of the step overwrites this rank's rows
            # the backward's scratch lives as long as the stepper and stays clean from call to call (no per-view clear);
            # dropped while a call is in flight so that an exception cannot leave a dirty one behind
            scratch = self._clean_scratch if self._clean_scratch is not None else C.clean_backward_scratch(P, dev)
            self._clean_scratch = None
            # the densification statistics of the view are updated on the way (the radii part on the side stream, the
            # gradient norms inside the per-Gaussian backward): no dense dL/dmeans2D, no separate pass over P at the end
            color, radii, img = be.view_fwd_bwd_packed(cam, dL, lv, base[k] + 64, base[k], None, after_forward,
                                                       clean_scratch=scratch, status_stream=side,
                                                       stats=(self.xyz_gradient_accum, self.denom, self.max_radii2D))
            self._clean_scratch = scratch
            held.append((radii, img))                # read on the side stream (radii: statistics, img: its status words)
            self.last_radii = radii
        main.wait_event(box["cleared"])
        # this rank's own rows go in first, before the barrier (i.e. while the slower peers are still rendering):
        # the first view is WRITTEN into the freshly cleared gradients (plain stores), further local views are added
        C.accumulate_packed(lv.means3D, [base[0] + 64], [base[0]], None, M, self.sh_degree, dsh, geom, store=True, marks=marks)
        if nv > 1:
            C.accumulate_packed(lv.means3D, [b + 64 for b in base[1:]], base[1:], None, M, self.sh_degree, dsh, geom, marks=marks)
        if G > 1:
            pk["hd"].barrier(channel=0)                 # every rank's rows (and headers) are in place
            # the peers' rows, read over NVLink; each rank walks its peers in a different rotation, so that at any
            # moment the ranks read from different peers
            me = dist.get_rank(self.group)
            order = [(me + j) % G for j in range(1, G)]
            bases = [pk["ptrs"][r] + 4 * k * stride for r in order for k in range(nv)]
            for lo in range(0, len(bases), 16):         # B200SPLAT_MAX_VIEWS sources per launch
                hi = min(lo + 16, len(bases))
                C.accumulate_packed(lv.means3D, [b + 64 for b in bases[lo:hi]], bases[lo:hi], None, M, self.sh_degree, dsh, geom, marks=marks)
            # (the barrier that lets the ranks overwrite their rows again sits at the START of the next step, beside
            # its forward: see `rows_free` above)
        side_done = torch.cuda.Event(); side_done.record(side)
        main.wait_event(side_done)
        del held
        # the geometry gradients are strided views of the interleaved [P,12] rows
        self.leaves.rotations.grad = geom[:, 0:4]
        self.leaves.means3D.grad = geom[:, 4:7]
        self.leaves.scales.grad = geom[:, 7:10]
        self.leaves.opacities.grad = geom[:, 10:11].view_as(self.leaves.opacities) if self.leaves.opacities.dim() == 2 else geom[:, 10]
        self.leaves.shs.grad = dsh
        return color

    # -- factorized exchange --------------------------------------------------------------------------------
    def _setup_symm(self, P, nv, dev):
        """Symmetric (peer-mapped) buffers: flat gradients [11 P, padded] and masked colour gradients [nv, P, 3]."""
        if self._symm is not None:
            return self._symm if self._symm.get("nv") == nv else None
        if not self.peer_memory or self.world_size == 1 or dev.type != "cuda" or self._backend is not None:
            self._symm = {}
            return None
        try:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem
            group = self.group if self.group is not None else dist.group.WORLD
            pad = 1024 * self.world_size
            nflat = (11 * P + pad - 1) // pad * pad
            flat = symm_mem.empty(nflat, dtype=torch.float32, device=dev)
            hf = symm_mem.rendezvous(flat, group)
            dcol = symm_mem.empty(nv * P * 3, dtype=torch.float32, device=dev)
            hd = symm_mem.rendezvous(dcol, group)
            flat.zero_()
            if not getattr(hf, "multicast_ptr", 0):
                raise RuntimeError("no multicast mapping")
            self._symm = {"nv": nv, "flat": flat, "dcol": dcol.view(nv, P, 3), "hd": hd, "group_name": group.group_name,
                          "ptrs": [int(p) for p in hd.buffer_ptrs]}
        except Exception as e:          # no peer mappings on this system
            if self.peer_memory is True:
                raise
            self._symm = {"error": repr(e)}
            return None
        return self._symm

    def _flat_views(self, flat):
        P = self.leaves.means3D.shape[0]
        return {"dL_drotations": flat[0:4 * P].view(P, 4), "dL_dmeans3D": flat[4 * P:7 * P].view(P, 3),
                "dL_dscales": flat[7 * P:10 * P].view(P, 3), "dL_dopacity": flat[10 * P:11 * P].view(P, 1)}

    def _step_factorized(self, cams, dL_images, reduce, on_forward):
        import torch.distributed as dist
        be = self._backend if self._backend is not None else _CudaBackend(self)
        lv = GaussianParams(*[t.detach() for t in self.leaves])
        P, M = lv.means3D.shape[0], lv.shs.shape[1]
        dev = lv.means3D.device
        cuda = dev.type == "cuda"
        G = self.world_size
        exchange = reduce and G > 1
        nv = len(cams)
        sy = self._setup_symm(P, nv, dev) if exchange else None
        if sy:
            flat, dcols = sy["flat"], sy["dcol"]          # peer-mapped, persistent: valid until the next step
        else:
            flat = torch.empty(11 * P, dtype=torch.float32, device=dev)
            dcols = torch.empty(nv, P, 3, dtype=torch.float32, device=dev)
        outs = self._flat_views(flat)
        campos = torch.stack([c.camera_center.to(dev).float() for c in cams]).contiguous()
        box = {"symm": sy}
        if cuda and self._side is None:
            # collectives on a HIGH-priority stream (their few CTAs must get SM slots while a full-grid kernel is
            # running), the rebuild kernel on a normal-priority one (it must not starve them)
            self._side = (torch.cuda.Stream(device=dev, priority=-1), torch.cuda.Stream(device=dev))
        if sy:   # camera centres of all views: tiny, gathered at step start, off the critical path
            main, comm = torch.cuda.current_stream(dev), self._side[0]
            ev0 = torch.cuda.Event(); ev0.record(main)
            comm.wait_event(ev0)
            with torch.cuda.stream(comm):
                campos.record_stream(comm)
                box["all_campos"] = torch.empty(G * nv, 3, dtype=torch.float32, device=dev)
                dist.all_gather_into_tensor(box["all_campos"], campos, group=self.group)

        def start_sh_exchange():
            """Runs as soon as the LAST local view's masked colour gradient is enqueued, i.e. before its per-Gaussian
            backward: the all-gather and the local rebuild of dL/dsh go to a side stream and overlap that kernel."""
            if cuda:
                main = torch.cuda.current_stream(dev)
                comm, rebuild = self._side
                ready = torch.cuda.Event(); ready.record(main)
                comm.wait_event(ready)
                if box["symm"]:
                    # peer-memory path: barrier (every rank's colour gradients are written), then the rebuild kernel
                    # loads all G x nv views straight from the peers' memory over NVLink, then a second barrier
                    # (every rank has finished reading) before anyone may overwrite its buffer in the next step
                    sy_ = box["symm"]
                    with torch.cuda.stream(comm):
                        sy_["hd"].barrier(channel=0)
                        gathered = torch.cuda.Event(); gathered.record(comm)
                    rebuild.wait_event(gathered)
                    with torch.cuda.stream(rebuild):
                        box["all_campos"].record_stream(rebuild)
                        ptrs = [sy_["ptrs"][r] + v * P * 12 for r in range(G) for v in range(nv)]
                        box["dsh"] = be.sh_grad_from_views(lv.means3D, ptrs, box["all_campos"], M, self.sh_degree)
                        sy_["hd"].barrier(channel=1)
                        box["done"] = torch.cuda.Event(); box["done"].record(rebuild)
                    return
                with torch.cuda.stream(comm):
                    if exchange:
                        all_dcols = torch.empty(G * nv, P, 3, dtype=torch.float32, device=dev)
                        all_campos = torch.empty(G * nv, 3, dtype=torch.float32, device=dev)
                        dcols.record_stream(comm); campos.record_stream(comm)
                        dist.all_gather_into_tensor(all_campos, campos, group=self.group)
                        dist.all_gather_into_tensor(all_dcols, dcols, group=self.group)
                        all_dcols.record_stream(rebuild); all_campos.record_stream(rebuild)
                        views, cp = [all_dcols[v] for v in range(G * nv)], all_campos
                    else:
                        views, cp = [dcols[v] for v in range(nv)], campos
                        dcols.record_stream(rebuild); campos.record_stream(rebuild)
                    gathered = torch.cuda.Event(); gathered.record(comm)
                rebuild.wait_event(gathered)
                with torch.cuda.stream(rebuild):
                    box["dsh"] = be.sh_grad_from_views(lv.means3D, views, cp, M, self.sh_degree)
                    box["done"] = torch.cuda.Event(); box["done"].record(rebuild)
            else:
                if exchange:
                    all_dcols = torch.empty(G * nv, P, 3, dtype=torch.float32, device=dev)
                    all_campos = torch.empty(G * nv, 3, dtype=torch.float32, device=dev)
                    dist.all_gather_into_tensor(all_campos, campos, group=self.group)
                    dist.all_gather_into_tensor(all_dcols, dcols, group=self.group)
                    views, cp = [all_dcols[v] for v in range(G * nv)], all_campos
                else:
                    views, cp = [dcols[v] for v in range(nv)], campos
                box["dsh"] = be.sh_grad_from_views(lv.means3D, views, cp, M, self.sh_degree)

        color = None
        for k, (cam, dL) in enumerate(zip(cams, dL_images)):
            if k == 0:
                tgt = dict(outs, dL_dcolors=dcols[0])
            else:
                tmp = torch.empty_like(flat)
                tgt = dict(self._flat_views(tmp), dL_dcolors=dcols[k])
            color, radii, dmeans2D = be.view_fwd_bwd(cam, dL, lv, tgt, on_forward,
                                                     start_sh_exchange if k == nv - 1 else None)
            if k > 0:
                flat += tmp
            if k == nv - 1:
                flat_done = self._reduce_flat(flat, exchange, cuda, dev)   # before the statistics: off the critical path
            self._accumulate_stats(dmeans2D, radii)
            self.last_radii = radii
        if cuda:
            if flat_done is not None:
                torch.cuda.current_stream(dev).wait_event(flat_done)
            torch.cuda.current_stream(dev).wait_event(box["done"])
            box["dsh"].record_stream(torch.cuda.current_stream(dev))
        self.leaves.means3D.grad = outs["dL_dmeans3D"]
        self.leaves.shs.grad = box["dsh"]
        self.leaves.opacities.grad = outs["dL_dopacity"].view_as(self.leaves.opacities)
        self.leaves.scales.grad = outs["dL_dscales"]
        self.leaves.rotations.grad = outs["dL_drotations"]
        return color

    def _reduce_flat(self, flat, exchange, cuda, dev):
        import torch.distributed as dist
        if exchange:   # the remaining 11 f32 / Gaussian: one flat all-reduce, concurrent with the SH rebuild
            if cuda:
                main, comm = torch.cuda.current_stream(dev), self._side[0]
                ev = torch.cuda.Event(); ev.record(main)
                comm.wait_event(ev)
                with torch.cuda.stream(comm):
                    if self._symm and self._symm.get("flat") is flat:
                        torch.ops.symm_mem.multimem_all_reduce_(flat, "sum", self._symm["group_name"])   # in-switch (NVLS)
                    else:
                        flat.record_stream(comm)
                        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
                    ev2 = torch.cuda.Event(); ev2.record(comm)
                return ev2                    # the caller makes the main stream wait at the END of the step
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        return None

    def reset_densification_stats(self):
        """Zero the rank-local statistics: what the reference does after it has consumed them
        (/root/reference/scene/gaussian_model.py:391-393, densification_postfix)."""
        self.xyz_gradient_accum.zero_(); self.denom.zero_(); self.max_radii2D.zero_()

    def reduce_densification_stats(self):
        """-> (xyz_gradient_accum, denom, max_radii2D) over ALL ranks' views since the last reset.

        In every exchange mode the three attributes are RANK-LOCAL accumulators that `step()` only ever adds to
        (norm per view, then summed; `denom` counts visible views; `max_radii2D` is a running maximum), exactly as
        the reference accumulates them between densification events (/root/reference/train.py:127,130-133).  This
        call sum- / max-reduces COPIES, so it can be made at any time and any number of times; the caller consumes
        the returned tensors and then calls `reset_densification_stats()`.  Sums and maxima commute with the
        accumulation over steps, so the values equal a single process that rendered all the views."""
        out = (self.xyz_gradient_accum.clone(), self.denom.clone(), self.max_radii2D.clone())
        if self.world_size > 1:
            import torch.distributed as dist
            dist.all_reduce(out[0], op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(out[1], op=dist.ReduceOp.SUM, group=self.group)
            dist.all_reduce(out[2], op=dist.ReduceOp.MAX, group=self.group)
        return out

    def all_reduce_grads(self):
        import torch.distributed as dist
        works = []
        for t in self.leaves:
            if t.grad is None:
                t.grad = torch.zeros_like(t)
            works.append(dist.all_reduce(t.grad, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
        for w in works:
            w.wait()

    def grads(self) -> GaussianParams:
        return GaussianParams(*[t.grad for t in self.leaves])

    def grad_checksum(self) -> torch.Tensor:
        return self.leaves.means3D.grad.sum() + self.leaves.opacities.grad.sum()

    # -- measurement helpers (CUDA path only) ------------------------------------------------------------
    def launches_per_view(self=None) -> int:
        """Kernels of THIS library launched per view fwd+bwd (memset / copy nodes are driver work and are not counted):
        preprocess 1; depth order 4 radix passes x (count, column scan, scatter) = 12, offsets spine 1; duplicateWithKeys 1;
        tile sort 2 passes x 3 = 6; tile order 1; render forward 1; render backward, compaction of the touched Gaussians,
        per-Gaussian backward 3.  Dense step: + background zero-fill of the gradient outputs 1, densification statistics 1.
        Packed step: + radii statistics 1, sparse clear of the dense gradients 1, count gather + accumulate 2 for the
        rank's own rows and 2 more for its peers'."""
        n = 1 + 12 + 1 + 1 + 6 + 1 + 1 + 3
        if self is not None and getattr(self, "exchange", None) == "packed":
            n += 1 + 1 + 2 + (2 if self.world_size > 1 else 0)
        else:
            n += 2
        return n

    def stage_profile(self, cam, dL, iters: int = 5):
        """Per-stage device milliseconds (CUDA events on the launch stream, inside the library) averaged
        over `iters` fwd+bwd passes, plus workload statistics.  The timed passes run the backward the packed step runs
        (64-byte rows out, clean scratch: no clear in front of the render backward, no dense outputs cleared beside
        it); the warm-up pass runs the dense backward, from which the number of touched Gaussians is read."""
        import diff_gaussian_rasterization as dgr
        C = dgr._C
        lv = GaussianParams(*[t.detach() for t in self.leaves])
        P = lv.means3D.shape[0]
        acc, stats = {}, {}
        scratch = C.clean_backward_scratch(P, lv.means3D.device)
        rows = torch.zeros(16 * P + 16, dtype=torch.float32, device=lv.means3D.device)
        cap = None          # first pass: exact (waits for R); timed passes: that R with head room, no host wait
        for it in range(iters + 1):
            st = C.StageTimes()
            res = C.rasterize_gaussians(self.bg, lv.means3D, None, lv.opacities, lv.scales, lv.rotations,
                                        self.scale_modifier, None, cam.world_view_transform, cam.full_proj_transform,
                                        cam.tanfovx, cam.tanfovy, cam.image_height, cam.image_width, lv.shs,
                                        self.sh_degree, cam.camera_center, False, False, stage_times=st,
                                        exact=cap is None, bin_capacity=cap)
            if cap is None:
                cap = max(1, int(1.25 * res[0]) + 65536)
            R, color, radii, geom, binning, img, _ = res
            if it > 0:
                C.rasterize_gaussians_backward_packed(
                    self.bg, lv.means3D, radii, lv.scales, lv.rotations, self.scale_modifier, cam.world_view_transform,
                    cam.full_proj_transform, cam.tanfovx, cam.tanfovy, dL, lv.shs, self.sh_degree, cam.camera_center, geom, R,
                    binning, img, lv.opacities, rows.data_ptr() + 64, rows.data_ptr(), cam.image_height, cam.image_width,
                    stage_times=st, clean_scratch=scratch)
                for k, v in st.as_dict().items():
                    acc[k[:-3]] = acc.get(k[:-3], 0.0) + v / iters
                continue
            g = C.rasterize_gaussians_backward(self.bg, lv.means3D, radii, None, lv.scales, lv.rotations,
                                           self.scale_modifier, None, cam.world_view_transform,
                                           cam.full_proj_transform, cam.tanfovx, cam.tanfovy, dL, lv.shs, self.sh_degree,
                                           cam.camera_center, geom, R, binning, img, False, opacities=lv.opacities,
                                           image_height=cam.image_height, image_width=cam.image_width, stage_times=st)
            if it == 0:
                stats = {"visible": int((radii > 0).sum().item()), "R": C.forward_status(img, cam.image_height, cam.image_width)[0],
                         "touched": int((g[3].abs().sum(dim=1) > 0).sum().item())}
                stats.update(C.render_stats(lv.means3D.shape[0], cam.image_width, cam.image_height, R, geom, binning, img))
        return acc, stats


class GraphedStep:
    """One whole step of a ViewParallelStep -- forward, backward, exchange, statistics: some 30 kernel launches and a
    dozen tensor operations per view -- captured ONCE in a CUDA graph and replayed with one launch per step.

    The step is launch-latency sensitive: at 2 ms per view the Python thread has to queue a kernel every 50 us, and on
    a host shared by 8 ranks (and their NCCL / NVML threads) it does not always manage; the replayed graph does not
    care.  What makes the capture possible: the pair count stays on the device (b200splat.h "pair capacity"), the
    packed exchange has no collective (camera centres travel in the row headers), and the symmetric-memory barrier is
    a stateless kernel.

    Inputs live in static device buffers (`load()` copies a step's cameras and upstream gradients into them); the
    intrinsics, image size and number of views are those of the cameras given at capture.  Outputs: `color`,
    `stepper.grads()`, the stepper's densification statistics -- valid until the next replay.  `step()` checks, one step
    late and without synchronising, that no replayed forward overflowed its pair capacity; if one did, it re-captures
    with the capacity the policy has meanwhile raised and raises for the spoiled step."""

    def __init__(self, stepper: ViewParallelStep, cams: Sequence, dL_images: Sequence[torch.Tensor], warmup: int = 3):
        if stepper.exchange not in ("packed",):
            raise ValueError("GraphedStep captures the packed-row step (exchange='packed' or 'auto' settled on packed)")
        self.st = stepper
        dev = stepper.leaves.means3D.device
        self.dev = dev
        self.cam_buf = [torch.empty(35, dtype=torch.float32, device=dev) for _ in cams]
        self.dL_buf = [torch.empty_like(d, device=dev) for d in dL_images]
        self.cams = [c._replace(world_view_transform=b[:16].view(4, 4), full_proj_transform=b[16:32].view(4, 4), camera_center=b[32:35])
                     for c, b in zip(cams, self.cam_buf)]
        self._shape = (cams[0].image_height, cams[0].image_width)
        self.load(cams, dL_images)
        self._capture(warmup)

    def _capture(self, warmup):
        st, dev = self.st, self.dev
        auto, st._auto = st._auto, False                  # the mode is fixed for the life of the graph
        cur = torch.cuda.current_stream(dev)
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(cur)
        with torch.cuda.stream(s):                        # eager warm-up: library streams, symmetric buffers, capacity history
            for _ in range(max(2, warmup)):
                st.step(self.cams, self.dL_buf)
        cur.wait_stream(s)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        # captured on a HIGH-priority stream: the kernels of the step's main chain keep that priority in the graph, the side
        # stream's background work (sparse clear of the dense gradients, radii statistics) fills in around them
        cap = torch.cuda.Stream(device=dev, priority=-1)
        with torch.cuda.graph(self.graph, stream=cap, capture_error_mode="thread_local"):
            self.color = st.step(self.cams, self.dL_buf)
        st._auto = auto

    def load(self, cams, dL_images=None):
        """Copy a step's cameras (and, if given, upstream gradients) into the static inputs; stream-ordered."""
        for b, c in zip(self.cam_buf, cams):       # one kernel per camera, not three copy-engine round trips
            torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1), c.camera_center.reshape(-1)], out=b)
        if dL_images is not None:
            for b, d in zip(self.dL_buf, dL_images):
                if b.data_ptr() != d.data_ptr():
                    b.copy_(d, non_blocking=True)

    def step(self, cams=None, dL_images=None) -> torch.Tensor:
        import diff_gaussian_rasterization as dgr
        status = dgr._C.capacity_status(self.dev, self.st.leaves.means3D.shape[0], *self._shape)
        if status is not None and status[2]:
            dgr._C.reset_capacity_policy()
            self._capture(2)                              # the exact first warm-up view sizes the new capacity
            raise RuntimeError(f"b200splat: a replayed forward produced {status[0]} (tile, Gaussian) pairs, more than the {status[1]} the "
                               "captured step was sized for; the step has been re-captured, re-run the spoiled step")
        if cams is not None:
            self.load(cams, dL_images)
        self.graph.replay()
        return self.color
