#!/usr/bin/env python
"""bench.py -- headline benchmark of the rasterizer hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl mine|reference]

A "step" is one mapping iteration of the hot path on one keyframe view: rasterizer forward +
backward (rasterize_gaussians + rasterize_gaussians_backward through the C ABI) on the headline
workload: 1M Gaussians, 1280x1024, SH degree 3 (M=16), 16-d feature. With N GPUs every rank
rasterizes its own view per step against replicated Gaussians and the parameter gradients are
summed with one NCCL all-reduce (weak scaling); value = view-iterations/s over all ranks.

--impl reference times the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built by
oracle/build_ref.sh for sm_100a) on the same workload on GPU 0 (the reference has no CPU path and
no multi-GPU path); if oracle/_ref is absent it falls back to timing the CPU oracle port.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "segment-any-gaussian-splatting-slam_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "mapping_iters_per_sec_fwd_bwd"
UNIT = "iters/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="mine", choices=["mine", "reference"])
    ap.add_argument("--P", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1280)
    ap.add_argument("--height", type=int, default=1024)
    ap.add_argument("--F", type=int, default=16)
    ap.add_argument("--deg", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--nccl-exchange", action="store_true",
                    help="N > 1, sparse exchange: NCCL all-gather + all-reduce instead of the peer-memory kernels")
    ap.add_argument("--no-exchange-overlap", action="store_true",
                    help="N > 1, sparse exchange: run the flag all-gather + compaction after the per-Gaussian backward "
                         "instead of beside it")
    ap.add_argument("--dense-allreduce", action="store_true",
                    help="N > 1: sum the full 4(11+3M+F) B/Gaussian bucket instead of the sparse (touched-rows) exchange")
    ap.add_argument("--distinct-views", action="store_true", help="rank r renders keyframe pose r (unequal work)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, enabled=True):
        self.index, self.rows, self.proc, self.first, self.enabled = index, [], None, 0, enabled

    def mark(self):
        """Call right before the timed region: only samples taken from here on are reported."""
        self.first = len(self.rows)

    def start(self):
        if os.environ.get("GSR_BENCH_NO_CLOCKS") or not self.enabled:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
            # nvidia-smi needs up to seconds before its first line on a multi-GPU box: wait for it here (untimed), so
            # that the sampler is running during the timed region
            t0 = time.time()
            while not self.rows and time.time() - t0 < 10.0 and self.proc.poll() is None:
                time.sleep(0.02)
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        rows = self.rows[self.first:] or self.rows
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def make_workload(args, view, device, grad_seed=99):
    """Seeded synthetic scene of the headline shape, tensors resident on `device`."""
    import numpy as np
    import torch
    from gsraster_b200 import scenes
    sc = scenes.make_scene(args.P, sh_degree=args.deg, F=args.F, seed=1234, W=args.width, H=args.height)
    cam = scenes.make_camera(args.width, args.height, view)
    grads = scenes.make_upstream_grads(args.width, args.height, args.F, seed=grad_seed)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
    d = dict(bg=t(sc.bg), means3D=t(sc.means3D), scales=t(sc.scales), rotations=t(sc.rotations),
             opacities=t(sc.opacities), shs=t(sc.shs), sh_objs=t(sc.sh_objs), viewmatrix=t(cam.viewmatrix),
             projmatrix=t(cam.projmatrix), campos=t(cam.campos))
    g = [t(x) for x in grads]
    return sc, cam, d, g


def stage_model(args, stats):
    """ALGORITHMIC bytes / flops per launch of each stage (restated in DESIGN.md).
    stats: P, Pvis, R, Npix, n_eval_fwd, n_blend, n_eval_bwd."""
    P, Pv, R, N = stats["P"], stats["Pvis"], stats["R"], stats["Npix"]
    M, F = (args.deg + 1) ** 2, args.F
    L1 = stats["l1_total"]
    return {
        "preprocess": ("hbm", Pv * (44 + 12 * M + 84) + (P - Pv) * 20),
        # 4 depth passes over P 8-byte items (histogram read + scatter read + write) + coverage grids (8 B/Gaussian)
        "depth_sort_offsets": ("hbm", P * (4 * 24 + 8)),
        # supertile lists: 2 sweeps over (depth key 8 + tiles_touched 4 + rect 4) B/Gaussian, 8 B/entry written;
        # (the per-tile pull, 4 B/instance, now lives in the forward blend and covers only the consumed prefix)
        "binning": ("hbm", P * 32 + L1 * 16),
        "render_fwd": ("fp32", 13 * stats["n_eval_fwd"] + (15 + 2 * (3 + F)) * stats["n_blend"]),
        "render_bwd": ("fp32", 13 * stats["n_eval_bwd"] + (100 + 7 * (3 + F)) * stats["n_blend"]),
        "preprocess_bwd": ("hbm", Pv * ((95 + 12 * M) + (64 + 12 * M)) + (P - Pv) * 4),
    }


def load_traffic():
    path = os.path.join(ROOT, "profiles", "traffic.json")
    return json.load(open(path)) if os.path.exists(path) else {}


def run_mine(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gsraster_b200 import capi, mapping
    dev = torch.device("cuda", local_rank)
    # Weak scaling needs equal work per rank. The SURVEY 8d scene sampler is centred on view 0, so other poses see a
    # different visible set (measured: 2.72 ms of kernels for view 0, 3.75 ms for view 1) and the all-reduce would
    # then time the synthetic scene's load imbalance. Default: every rank renders the view-0 pose with its own
    # upstream gradients; --distinct-views gives rank r the pose of view r.
    sc, cam, d, g = make_workload(args, view=rank if args.distinct_views else 0, device=dev, grad_seed=99 + rank)
    P, W, H, F, deg = args.P, args.width, args.height, args.F, args.deg
    M = (deg + 1) ** 2
    e = torch.empty(0, device=dev)
    sparse = world > 1 and not args.dense_allreduce
    peer = sparse and not args.nccl_exchange
    bucket = mapping.GradBucket(P, M, F, dev, symmetric=peer)
    launches = [0]
    ex = mapping.SparseGradExchange(bucket, peer_memory=peer) if sparse else None
    if sparse:
        ex.arm(overlap=not args.no_exchange_overlap)   # touched flags + blend-done event registered with the library

    def fwd(dd):
        return capi.rasterize_gaussians(dd["bg"], dd["means3D"], e, dd["opacities"], dd["scales"], dd["rotations"],
                                        1.0, e, dd["viewmatrix"], dd["projmatrix"], cam.tanfovx, cam.tanfovy, H, W,
                                        dd["shs"], dd["sh_objs"], deg, dd["campos"], False, False)

    def bwd(dd, res, gd, gc, go):
        R, depth, color, objects, radii, is_used, gb, bb, ib = res
        return capi.rasterize_gaussians_backward(dd["bg"], dd["means3D"], radii, e, dd["scales"], dd["rotations"], 1.0,
                                                 e, dd["viewmatrix"], dd["projmatrix"], cam.tanfovx, cam.tanfovy, gd,
                                                 gc, dd["shs"], go, dd["sh_objs"], deg, dd["campos"], gb, R, bb, ib,
                                                 False, out=bucket.views)

    def step():
        res = fwd(d)
        launches[0] += capi.last_launch_count()
        if sparse:
            ex.begin_step()
        bwd(d, res, g[0], g[1], g[2])
        launches[0] += capi.last_launch_count()
        if world > 1:
            if ar_events is not None:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
            if peer:
                ex.exchange()      # barrier, union of the peers' flags, rank-ordered row sums out of the peers' buckets,
                                   # barrier: all queued, nothing to read back (capacity = P; consumers take device count)
            elif sparse:
                ex.exchange()      # flags all-gather, union scan, row gather, SUM-reduce of the touched rows: all queued
                ex.finish()        # count read-back (dense fallback if the capacity guess was too small)
            else:
                dist.all_reduce(bucket.flat)
            if ar_events is not None:
                b.record()
                ar_events.append((a, b))
        return res

    ar_events = None

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        res = step()
    sync_all()
    # workload statistics for the roofline model (outside the timed region)
    R = res[0]
    l1_total = capi.last_forward_stats()["l1_total"]
    st = capi.export_state(P, deg, M, F, W, H, R, res[6], res[7], res[8])
    cp = capi.count_pairs(P, deg, M, F, W, H, R, res[6], res[7], res[8])
    stats = dict(P=P, Pvis=int((res[4] > 0).sum()), R=int(R), Npix=W * H, n_eval_fwd=cp["evaluated"],
                 n_blend=cp["blended"], n_eval_bwd=int(st["n_contrib"].to(torch.int64).sum()),
                 warp_blended=cp["warp_blended"], l1_total=l1_total)
    del st
    tm = torch.zeros(P, dtype=torch.uint8, device=dev)
    capi.set_touched_mask(tm)
    bwd(d, res, g[0], g[1], g[2])
    capi.set_touched_mask(ex.touched if sparse else None)
    touched_rows = int(tm.sum())
    del tm
    clocks = ClockSampler(local_rank, enabled=(rank == 0))  # rank 0 prints the line; 8 concurrent nvidia-smi slow each other
    clocks.start()  # waits (untimed) for nvidia-smi's first sample, so the sampler is live during the timed region
    sync_all()      # the other ranks wait here, not inside a device-side barrier
    for _ in range(3):  # settle the caching allocator again after the statistics pass (untimed, in addition to W)
        step()
    sync_all()

    # ---- timed region 1: inputs resident in HBM, EXACTLY K steps, nothing but the step on the stream ----
    launches[0] = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    clocks.mark()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    sync_all()
    ms = ev0.elapsed_time(ev1)
    n_launch = launches[0]

    # ---- region 1b: the same K steps again with every stage bracketed by CUDA events (roofline figures). The 12
    # event records per step serialise the stream and lengthen the num_rendered bubble (measured: +0.4 ms/step),
    # so this pass is reported separately (ms_per_step_instrumented) and never used for `value`. ----
    instrument = not os.environ.get("GSR_BENCH_NO_STAGE_TIMING")
    capi.set_stage_timing(instrument)
    capi.get_stage_times()
    capi.get_stage_gaps()
    ar_events = [] if (world > 1 and instrument) else None
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev2.record()
    for _ in range(args.steps if instrument else 0):
        step()
    ev3.record()
    sync_all()
    ms_instr = ev2.elapsed_time(ev3)
    stage_ms = capi.get_stage_times()
    gap_ms = capi.get_stage_gaps()
    capi.set_stage_timing(False)
    allreduce_ms = sum(a.elapsed_time(b) for a, b in ar_events) / len(ar_events) if ar_events else 1e-9
    ar_events = None
    if peer and ex._pending is not None:
        ex.finish()   # outside every timed region: the union size of the last step, for the report

    # ---- timed region 2: end to end with host buffers ----
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, cam, d, g, fwd, bwd, bucket, world, dev, sync_all, ex)
    clk = clocks.stop()
    rank_ms = [sum(stage_ms.values())]
    if world > 1:  # per-rank kernel time (sum of the six stages): shows the skew the all-reduce has to absorb
        allr = [torch.zeros(1, device=dev, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(allr, torch.tensor([rank_ms[0]], device=dev, dtype=torch.float64))
        rank_ms = [round(float(x), 3) for x in allr]
        t = torch.tensor([ms, e2e["ms"] if e2e else 0.0], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        if e2e:
            e2e["ms"] = float(t[1])
    return dict(ms=ms, ms_instr=ms_instr, gap_ms=gap_ms, stage_ms=stage_ms, stats=stats, launches=n_launch, e2e=e2e, clocks=clk,
                bucket_bytes=bucket.nbytes, allreduce_ms=allreduce_ms, rank_ms=rank_ms,
                exchange_rows=(ex.last_rows if sparse else P), touched_rows=touched_rows)


def run_e2e(args, cam, d, g, fwd, bwd, bucket, world, dev, sync_all, ex=None):
    """Same step through the public API with HOST buffers: per step the keyframe (view matrix,
    projection matrix, camera centre, ground-truth RGB image) is copied from pinned host memory, an
    the reference's masked L1 photometric loss (gsr_photometric_loss, lambda_dssim = 0) on the rendered colour produces dL/dcolour on the device, and the loss
    scalar is read back. Gaussian parameters stay resident (they are the optimiser's state)."""
    import torch
    import torch.distributed as dist
    W, H = args.width, args.height
    pin = lambda t: t.detach().cpu().pin_memory()
    h_view, h_proj, h_campos = pin(d["viewmatrix"]), pin(d["projmatrix"]), pin(d["campos"])
    torch.manual_seed(7)
    h_gt = torch.rand(3, H, W).pin_memory()
    dd = dict(d)
    dd["viewmatrix"], dd["projmatrix"], dd["campos"] = (torch.empty_like(d["viewmatrix"]),
                                                         torch.empty_like(d["projmatrix"]),
                                                         torch.empty_like(d["campos"]))
    gt = [torch.empty(3, H, W, device=dev) for _ in range(2)]  # double-buffered: the copy of step i overlaps step i's forward
    h_loss = [torch.zeros(1).pin_memory() for _ in range(2)]
    h2d = sum(t.numel() * 4 for t in (h_view, h_proj, h_campos, h_gt))
    from gsraster_b200 import losses
    ploss = losses.PhotometricLoss(H, W, 0.0, device=dev)  # lambda_dssim = 0: masked L1 (the SSIM kernels are measured in loss_n2)
    loss3 = [torch.empty(3, device=dev) for _ in range(2)]
    grad_buf = torch.empty(3, H, W, device=dev)
    copy_stream = torch.cuda.Stream(device=dev)
    ev_gt = [torch.cuda.Event() for _ in range(2)]
    ev_free = [torch.cuda.Event() for _ in range(2)]
    ev_loss = [torch.cuda.Event() for _ in range(2)]
    main = torch.cuda.current_stream()
    for e_ in ev_free:
        e_.record(main)
    state = dict(i=0, last=0.0)

    def step():
        i = state["i"]
        k = i & 1
        # ground-truth image: pinned host -> device on the copy stream, concurrent with this step's forward
        copy_stream.wait_event(ev_free[k])
        with torch.cuda.stream(copy_stream):
            gt[k].copy_(h_gt, non_blocking=True)
            ev_gt[k].record(copy_stream)
        dd["viewmatrix"].copy_(h_view, non_blocking=True)
        dd["projmatrix"].copy_(h_proj, non_blocking=True)
        dd["campos"].copy_(h_campos, non_blocking=True)
        res = fwd(dd)
        main.wait_event(ev_gt[k])
        # masked L1 photometric loss (loss_utils.py:17-20), value and dL/dcolour in one fused pass
        out3, grad_c = ploss(res[2], gt[k], out=loss3[k], grad_out=grad_buf)
        loss = out3[0:1]
        ev_free[k].record(main)
        if ex is not None:
            ex.begin_step()
        bwd(dd, res, g[0], grad_c, g[2])
        if world > 1:
            if ex is not None:
                ex.exchange()
                if not ex.peer:
                    ex.finish()
            else:
                dist.all_reduce(bucket.flat)
        h_loss[k].copy_(loss.reshape(1), non_blocking=True)  # 4-byte D2H of the step's result
        ev_loss[k].record(main)
        if i > 0:  # read the previous step's loss: keeps one step in flight instead of draining the GPU
            ev_loss[k ^ 1].synchronize()
            state["last"] = float(h_loss[k ^ 1][0])
        state["i"] = i + 1
        return state["last"]

    for _ in range(3):
        step()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev_loss[(state["i"] - 1) & 1].synchronize()
    loss = float(h_loss[(state["i"] - 1) & 1][0])
    ev1.record()
    sync_all()
    return dict(ms=ev0.elapsed_time(ev1), h2d=h2d, d2h=4, loss=loss)


def optimizer_leg(args, touched_rows, hbm_gbs):
    """SURVEY 8f N1 (next row, measured beside the headline): fused activation + Adam on the headline shapes.
    dense = torch.optim.Adam semantics over all P rows; sparse = only the rows a view's backward pass reached."""
    import torch
    from gsraster_b200 import optim
    P, F, M = args.P, args.F, (args.deg + 1) ** 2
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev).manual_seed(3)
    r = lambda *s: torch.randn(*s, generator=g, device=dev)
    raw = dict(xyz=r(P, 3), f_dc=r(P, 1, 3), f_rest=r(P, M - 1, 3), opacity=r(P, 1), scaling=r(P, 3) - 3.0,
               rotation=r(P, 4), objects=r(P, 1, F))
    fa = optim.FusedGaussianAdam(raw)
    grads = dict(means3D=r(P, 3), sh=r(P, M, 3), objects=r(P, F), opacity=r(P, 1), scales=r(P, 3), rotations=r(P, 4))
    U = max(int(touched_rows), 1)
    idx = torch.sort(torch.randperm(P, device=dev)[:U])[0].to(torch.int32)
    cnt = torch.tensor([U], dtype=torch.int32, device=dev)

    def timed(fn, n=20):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    dense_ms = timed(lambda: fa.step(grads))
    sparse_ms = timed(lambda: fa.step(grads, idx=idx, count=cnt))
    nfl = 11 + 3 * M + F
    dense_bytes = P * 4 * (7 * nfl + 3 * M + 8)   # p, m, v read + written, gradient read, activated values written
    return dict(dense_ms=round(dense_ms, 4), dense_gbs=round(dense_bytes / dense_ms / 1e6, 1),
                dense_frac_of_hbm=round(dense_bytes / dense_ms / 1e6 / hbm_gbs, 3), sparse_rows=U,
                sparse_ms=round(sparse_ms, 4),
                note="fused activation + Adam (gsr_adam_step): dense = torch.optim.Adam over all P rows, "
                     "sparse = the rows one view's backward pass reached; not part of `value`")


def loss_leg(args, hbm_gbs):
    """SURVEY 8f N2 (next row): fused photometric loss (value + gradient) on the headline image size."""
    import torch
    from gsraster_b200 import losses
    H, W = args.height, args.width
    dev = torch.device("cuda", torch.cuda.current_device())
    gt = torch.rand(3, H, W, device=dev)
    img = (gt + 0.1 * torch.randn(3, H, W, device=dev)).clamp(0, 1)
    pl = losses.PhotometricLoss(H, W, 0.2, device=dev)
    out3, grad = torch.empty(3, device=dev), torch.empty_like(img)
    for _ in range(3):
        pl(img, gt, out=out3, grad_out=grad)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        pl(img, gt, out=out3, grad_out=grad)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    nbytes = 3 * H * W * 4 * (2 + 3 + 3 + 2 + 1)  # img, gt read; 3 maps written; 3 maps read; img, gt read; gradient written
    return dict(ms=round(ms, 4), gbs=round(nbytes / ms / 1e6, 1), frac_of_hbm=round(nbytes / ms / 1e6 / hbm_gbs, 3),
                note="gsr_photometric_loss: (1 - 0.2) L1 + 0.2 (1 - SSIM) with the reference's masks, value + dL/dimage; "
                     "part of the e2e leg, not of `value`")


def feature_loss_leg(args, hbm_gbs):
    """SURVEY 8f N2 (next row): fused classifier + cross-entropy + argmax on the headline feature image, with the torch
    module composition the reference runs (Conv2d 1x1 + CrossEntropyLoss + autograd) timed beside it."""
    import torch
    from gsraster_b200 import losses
    H, W, F, C = args.height, args.width, args.F, 100
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(7)
    classifier = torch.nn.Conv2d(F, C, kernel_size=1).to(dev)
    objects = torch.randn(F, H, W, device=dev)
    gt = torch.randint(0, C, (H, W), device=dev)
    fl = losses.FeatureLoss(classifier, device=dev)
    out2, grad = torch.empty(2, device=dev), torch.empty_like(objects)

    def timed(fn, n):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    ms = timed(lambda: fl(objects, gt, need_pred=True, out=out2, grad_out=grad), 20)
    crit = torch.nn.CrossEntropyLoss(reduction='none')
    xo = objects.clone().requires_grad_(True)

    def torch_modules():
        xo.grad = None
        logits = classifier(xo)
        torch.argmax(logits, dim=0)
        (crit(logits.unsqueeze(0), gt.unsqueeze(0)).squeeze().mean() / 4.605170186).backward()

    tms = timed(torch_modules, 5)
    flops = 2.0 * 3 * 2 * F * 104 * H * W  # two products, 3xTF32, classes padded to 104
    nbytes = H * W * (4 * F + 8 + 4 * F + 4)
    return dict(ms=round(ms, 4), torch_modules_ms=round(tms, 4), tensor_tflops=round(flops / ms / 1e9, 1),
                gbs=round(nbytes / ms / 1e6, 1), frac_of_hbm=round(nbytes / ms / 1e6 / hbm_gbs, 3),
                note="gsr_feature_loss: Conv2d(F,100,1) + cross-entropy / ln C + argmax, value + dL/dobjects, logits kept in "
                     "registers (mma.sync 3xTF32); torch_modules_ms = the same through torch.nn + autograd on this GPU; "
                     "not part of `value`")


def cls3d_leg(args):
    """SURVEY 8f N4 (next row): fused 3-D feature-consistency regulariser at the reference's call-site sizes (300 000-point
    subset, 1 000 samples, k = 5, C = 100) on the headline map, with the torch composition the reference runs (classifier
    + softmax over ALL Gaussians, cdist, topk, KL, autograd) timed beside it."""
    import torch
    from gsraster_b200 import losses
    P, F, C = args.P, args.F, 100
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(11)
    classifier = torch.nn.Conv2d(F, C, kernel_size=1).to(dev)
    xyz = torch.randn(P, 3, device=dev) * 4
    objects_dc = torch.randn(P, 1, F, device=dev)
    fn = losses.Cls3DLoss(classifier, device=dev)
    subset, sample = losses.Cls3DLoss.draw_indices(P, 300000, 1000, device=dev, generator=torch.Generator().manual_seed(2))
    grad, out1 = torch.zeros(P, F, device=dev), torch.empty(1, device=dev)

    def timed(f, n):
        for _ in range(2):
            f()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    ms = timed(lambda: fn(xyz, objects_dc.view(P, F), subset, sample, grad_out=grad, out=out1), 10)
    xo = objects_dc.clone().requires_grad_(True)
    sub64 = None if subset is None else subset.long()
    smp64 = sample.long()

    def torch_composition():   # scripts/gaussian_splatting.py:812-814 + utils/loss_utils.py:86-112 with the draws fixed
        xo.grad = None
        prob = torch.softmax(classifier(xo.permute(2, 0, 1)), dim=0).squeeze().permute(1, 0)
        feats, preds = (xyz, prob) if sub64 is None else (xyz[sub64], prob[sub64])
        sp = preds[smp64]
        nb = torch.cdist(feats[smp64], feats).topk(5, largest=False).indices
        kl = sp.unsqueeze(1) * (torch.log(sp.unsqueeze(1) + 1e-10) - torch.log(preds[nb] + 1e-10))
        (2.0 * kl.sum(dim=-1).mean() / C).backward()

    tms = timed(torch_composition, 3)
    return dict(ms=round(ms, 4), torch_composition_ms=round(tms, 4), subset=int(300000 if subset is not None else P), samples=1000,
                note="gsr_cls3d_loss: kNN (k=5) of 1 000 samples in a 300 000-point subset + classifier/softmax/KL of the "
                     "1 + k points of each sample + gradient scattered to dL/dobjects; torch_composition_ms = the "
                     "reference's lines in torch on this GPU; every 5th mapping iteration, not part of `value`")


def store_leg(args):
    """SURVEY 8f N3 (next row): map growth and pruning on the capacity-managed store against the reference's torch
    operations (torch.cat of the 21 parameter / Adam-moment tensors per frame; boolean-mask indexing of them per prune)."""
    import torch
    from gsraster_b200 import store
    from gsraster_b200.optim import GROUPS
    P, M, F, n_new = args.P, (args.deg + 1) ** 2, args.F, 30000
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev).manual_seed(3)
    shapes = store.row_shapes(M, F)
    rows = lambda n: {k: torch.randn((n,) + s, generator=g, device=dev) for k, s in shapes.items()}
    base, new = rows(P), rows(n_new)
    base["opacity"] *= 3.0

    def timed(f, n):
        f()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(n):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    st = store.GaussianStore(P + 8 * n_new, M, F, device=dev)
    st.append(base)

    def grow():
        st.count = P
        st.append(new)

    grow_ms = timed(grow, 5)
    st.count = P

    f0 = st.front

    def prune():
        st.front, st.count = f0, P      # compaction is out of place: the original rows are still in set f0
        st.prune_large_and_transparent(0.3, None)

    prune_ms = timed(prune, 3)
    kept = st.count
    st.count = 0
    del st
    ref = [{k: v.clone() for k, v in base.items()} for _ in range(3)]

    def torch_grow():   # cat_tensors_to_optimizer: parameters + zero moments
        return [{k: torch.cat((r[k], new[k] if i == 0 else torch.zeros_like(new[k])), dim=0) for k in GROUPS}
                for i, r in enumerate(ref)]

    def torch_prune():  # prune_large_and_transparent + _prune_optimizer
        valid = ~(torch.sigmoid(ref[0]["opacity"]) < 0.3).squeeze()
        return [{k: r[k][valid] for k in GROUPS} for r in ref]

    tg, tp = timed(torch_grow, 3), timed(torch_prune, 3)
    return dict(grow_ms=round(grow_ms, 4), torch_cat_ms=round(tg, 4), prune_ms=round(prune_ms, 4), torch_mask_index_ms=round(tp, 4),
                new_rows=n_new, kept_rows=kept,
                note="GaussianStore.append (new rows written behind the count) / prune_large_and_transparent (flags, "
                     "compaction list, 3 row-gather launches, count read-back) at P rows vs torch.cat / boolean-mask "
                     "indexing of the 21 parameter and Adam-moment tensors on this GPU; not part of `value`")


def cpu_baseline(args, threads=None):
    """CPU oracle port (oracle/oracle.cpp, OpenMP) timed on the host cores on a bounded sample."""
    import numpy as np
    from gsraster_b200 import scenes
    from oracle import pyoracle as po
    po.build()
    cores = po.num_threads()
    # bounded sample: the full camera/resolution with P scaled so the sample stays in the 10-30 s range
    Ps = args.P if cores >= 32 else min(args.P, 250_000)
    sc = scenes.make_scene(Ps, sh_degree=args.deg, F=args.F, seed=1234, W=args.width, H=args.height)
    cam = scenes.make_camera(args.width, args.height, 0)
    grads = scenes.make_upstream_grads(args.width, args.height, args.F)
    pr = po.Problem(sc, cam)
    t0 = time.time()
    st = po.forward(pr)
    po.backward(pr, st, *grads)
    dt = time.time() - t0
    sample = "1 fwd+bwd iteration, %d Gaussians (of %d), %dx%d, deg %d, F=%d, R=%d" % (
        Ps, args.P, args.width, args.height, args.deg, args.F, st["R"])
    return dict(value=1.0 / dt, unit=UNIT, cores=cores, kind="port", sample=sample, seconds=dt)


def run_reference(args):
    """Reference arm: unmodified reference CUDA rasterizer (oracle/_ref) on GPU 0, same workload."""
    import torch
    from oracle import refcuda
    cfg = dict(workload="headline: 1 view fwd+bwd, %d Gaussians, %dx%d, SH deg %d, F=%d" % (
        args.P, args.width, args.height, args.deg, args.F), l2="inputs (>=300 MB) larger than the 126 MB L2")
    if not (torch.cuda.is_available() and refcuda.available(args.F)):
        cb = cpu_baseline(args)
        ms = 1e3 * cb["seconds"]
        return dict(metric=METRIC, value=cb["value"], unit=UNIT, n_gpus=0, steps=1, warmup=0, ms_per_step=ms,
                    higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                    config=cfg, impl="reference", cpu_baseline={k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
                    e2e=dict(value=cb["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    note="oracle/_ref (reference CUDA build) not available: CPU oracle port timed instead")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sc, cam, d, g = make_workload(args, view=0, device=dev)
    ref = refcuda.RefRasterizer(args.F)
    W, H = args.width, args.height

    def step():
        R, depth, color, objects, radii, is_used, bufs = ref.forward(d, W, H, cam.tanfovx, cam.tanfovy, args.deg)
        ref.backward(d, W, H, cam.tanfovx, cam.tanfovy, args.deg, R, radii, bufs, g[0], g[1], g[2])
        return R

    for _ in range(max(args.warmup, 3)):
        R = step()
    torch.cuda.synchronize()
    clocks = ClockSampler(0)
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.mark()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop()
    value = args.steps / (ms * 1e-3)
    return dict(metric=METRIC, value=value, unit=UNIT, n_gpus=1, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ms / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=dict(cfg, R=int(R)), impl="reference", clocks=clk,
                cpu_baseline=dict(value=value, unit=UNIT, cores=0, kind="reference",
                                  sample="unmodified reference CUDA rasterizer built for sm_100a (oracle/_ref), "
                                         "run on GPU 0: the reference has no CPU implementation of this path"),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args)), flush=True)
        return 0
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its version / INFO lines to stdout by default; stdout of rank 0 carries exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    r = run_mine(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return 0
    pk = peaks()
    K = args.steps
    ms = r["ms"]
    value = world * K / (ms * 1e-3)
    # per-stage roofline (CUDA events around every stage in the instrumented pass, per launch = per step)
    model = stage_model(args, r["stats"])
    fp32_peak = 148 * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12  # TFLOP/s at clocks.max.sm, non-tensor FMA pipes
    traffic = load_traffic()
    stages = {}
    for name, (bound, work) in model.items():
        t = max(r["stage_ms"][name] / K * 1e-3, 1e-12)
        if bound == "hbm":
            ach, peak, unit = work / t / 1e9, pk["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = work / t / 1e12, fp32_peak, "TFLOP/s"
        stages[name] = dict(ms=round(t * 1e3, 4), bound=bound, achieved=round(ach, 3), peak=round(peak, 1), unit=unit,
                            frac=round(ach / peak, 4), work=int(work), traffic=traffic.get(name))
    top = max(stages, key=lambda k: stages[k]["ms"])
    roof = dict(kernel=top, **{k: stages[top][k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")},
                peak_source=("MEASURED_PEAKS.json hbm_gbs" if stages[top]["bound"] == "hbm" else
                             "148 SM x 128 FMA lanes x 2 x clocks.max.sm (no fp32 figure in MEASURED_PEAKS.json)") +
                            " (%s)" % pk["source"])
    out = dict(metric=METRIC, value=round(value, 4), unit=UNIT, n_gpus=world, steps=K, warmup=max(args.warmup, 3),
               ms_per_step=round(ms / K, 4), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
               data="synthetic",
               config=dict(workload="headline: 1 view fwd+bwd per GPU per step, %d Gaussians, %dx%d, SH deg %d (M=%d), "
                                    "F=%d feature channels" % (args.P, args.width, args.height, args.deg,
                                                               (args.deg + 1) ** 2, args.F),
                           parallelism="view-data-parallel x%d (%s), replicated Gaussians, %s" % (
                               world, "pose of view r on rank r" if args.distinct_views else
                               "same pose, per-rank upstream gradients: equal work per rank",
                               "1 NCCL all-reduce of the dense %d MB bucket/step" % (r["bucket_bytes"] // 1000000)
                               if args.dense_allreduce else
                               ("sparse gradient exchange: NCCL all-gather of the touched flags + SUM-reduce of the " if args.nccl_exchange else
                                "sparse gradient exchange over peer memory (no collective library call: barrier, union of the "
                                "peers' touched flags read in place, rank-ordered sum of the union's rows out of the peers' "
                                "buckets over NVLink): ") +
                               "%d touched rows (of %d) = %d MB/step" % (
                                   r["exchange_rows"], args.P,
                                   r["exchange_rows"] * (r["bucket_bytes"] // args.P) // 1000000)) if world > 1 else "single GPU",
                           l2="inputs (>=300 MB of Gaussian parameters) larger than the 126 MB L2; no flush needed",
                           **{k: r["stats"][k] for k in ("Pvis", "R", "l1_total", "n_eval_fwd", "n_blend", "warp_blended", "n_eval_bwd")}),
               gpu_launches=r["launches"], clocks=r["clocks"], roofline=roof, stages=stages,
               ms_per_step_instrumented=round(r["ms_instr"] / K, 4),
               idle_before_stage_ms={k: round(v / K, 4) for k, v in r["gap_ms"].items()})
    if world > 1:
        xbytes = r["exchange_rows"] * (r["bucket_bytes"] // args.P)
        out["allreduce"] = dict(ms=round(r["allreduce_ms"], 4), bytes=xbytes, dense_bytes=r["bucket_bytes"],
                                rows=r["exchange_rows"], mode="dense" if args.dense_allreduce else "sparse",
                                busbw_gbs=round(2 * (world - 1) / world * xbytes / r["allreduce_ms"] / 1e6, 1),
                                rank_kernel_ms_per_step=[round(x / K, 4) for x in r["rank_ms"]])
    if r["e2e"]:
        e = r["e2e"]
        out["e2e"] = dict(value=round(world * K / (e["ms"] * 1e-3), 4), unit=UNIT, h2d_bytes_per_step=e["h2d"],
                          d2h_bytes_per_step=e["d2h"], ms_per_step=round(e["ms"] / K, 4))
    if world == 1:
        try:
            out["optimizer_n1"] = optimizer_leg(args, r["touched_rows"], pk["hbm_gbs"])
        except Exception as ex:  # the leg is informative only; the headline line must still be printed
            out["optimizer_n1"] = dict(error=repr(ex)[:200])
        try:
            out["loss_n2"] = loss_leg(args, pk["hbm_gbs"])
        except Exception as ex:
            out["loss_n2"] = dict(error=repr(ex)[:200])
        try:
            out["feature_loss_n2"] = feature_loss_leg(args, pk["hbm_gbs"])
        except Exception as ex:
            out["feature_loss_n2"] = dict(error=repr(ex)[:200])
        try:
            out["cls3d_n4"] = cls3d_leg(args)
        except Exception as ex:
            out["cls3d_n4"] = dict(error=repr(ex)[:200])
        try:
            out["store_n3"] = store_leg(args)
        except Exception as ex:
            out["store_n3"] = dict(error=repr(ex)[:200])
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline(args)
        out["cpu_baseline"] = {k: (round(cb[k], 5) if k == "value" else cb[k]) for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(out), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
