#!/usr/bin/env python3
"""Benchmark of the render hot path (CameraRays -> FullPathTracer) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cornell]

A "step" is one frame = one call of the renderer's render() = `--samples-per-frame` (default 16) sample passes of the whole
1080p frame per GPU, the reference's samples_per_frame setting (GPURenderer::render, src/Renderer/GPURenderer.cpp:408-464; per pass
two launch points: CameraRays, FullPathTracer; here a frame's passes run as fused sample batches where the options allow it; the path tracer
runs as a wavefront of per-bounce kernels: ray-queue traversal, surface, queue sort, shading phases).
Metric (BASELINE.json): Mrays/s = BVH queries (trace_ray iterations + shadow rays + shadow-light rays)
per second, whole job. Rays are counted by the counting build of the SAME kernels on the SAME seeds in an untimed
pass; the timed passes run the timing build. N > 1: every rank renders its own sample passes of the full frame
(sample-pass sharding, weak scaling) and the accumulation buffers are summed with one NCCL all-reduce per K steps
inside the timed region. `--impl reference` times the reference's own CPU kernels (oracle/_ref) on the host cores.
Prints ONE JSON line."""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons with nvidia-smi DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu_index, self.samples, self.stop_flag = gpu_index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.samples[0][1]) if self.samples[0][1].isdigit() else None, "reasons": reasons}


def measured_hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json, copy bandwidth)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(workload, width, height, spf):
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the k_trace_queue launches of ONE step, from the committed
    `ncu --set full` capture of `bench.py --profile` (profiles/traffic.json, written by tools/ncu_traffic.py); null if this
    workload / size was not captured."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(path):
        return None
    entry = json.load(open(path)).get(f"{workload}_{width}x{height}_spf{spf}")
    return (entry["trace_queue_dram_bytes_per_step"], entry.get("ncu_time_weighted")) if entry else (None, None)


class oracle_prints_to_stderr:
    """The reference's CPU kernels print diagnostics ('NaN at [x, y], sample n') on stdout from C++; this process must print ONE JSON line
    there. While the oracle runs, file descriptor 1 points at stderr."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *exc):
        os.dup2(self.saved, 1)
        os.close(self.saved)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def bounded_oracle(ref, wl, pt, cfg, seconds_per_pass=4.0):
    """The reference's CPU renderer on a BOUNDED sample of the workload: the whole frame when one sample pass takes under
    `seconds_per_pass`, else a quarter-size frame of the same scene and camera (1/16 of the pixels; rays per pixel, hence
    Mrays/s, do not depend on it). The OpenMP team is set to every host core explicitly: torch.distributed.run exports
    OMP_NUM_THREADS=1, which the reference's `#pragma omp parallel for` (CPURenderer.cpp:378-389) would otherwise inherit."""
    r = wl.setup_oracle(ref, pt.structs, cfg)
    cores = host_cores()
    r.lib.ref_set_num_threads(C_int(cores))
    W, H = cfg["width"], cfg["height"]
    r.resize(max(W // 4, 8), max(H // 4, 8))
    r.render(1)
    t0 = time.time()
    r.render(1)
    t_small = time.time() - t0
    small = t_small * 16.0 > seconds_per_pass
    if not small:
        r.resize(W, H)
        r.render(1)
    frame = f"{max(W // 4, 8)}x{max(H // 4, 8)} frame (1/16 of the pixels, same scene and camera)" if small else "whole frame"
    return r, r.lib.ref_num_threads(), (t_small if small else t_small * 16.0), frame, (16.0 if small else 1.0)


def C_int(v):
    import ctypes
    return ctypes.c_int(v)


def workload_config(cfg, args, extra=None):
    """`config` of the JSON line: the same keys on the GPU arm and on the reference arm."""
    out = {"workload": cfg["name"], "width": cfg["width"], "height": cfg["height"], "nb_bounces": cfg["settings"].get("nb_bounces", 8),
           "kernel_options": cfg["options"], "samples_per_frame": cfg["settings"].get("samples_per_frame", 1),
           "step": "one frame = samples_per_frame sample passes (1 spp each) of the whole image"}
    out.update(extra or {})
    return out


def run_reference(args, cfg, pt):
    """Reference arm: the reference's own CPU implementation of the path (oracle/_ref = its unmodified kernels) on every host
    core, each step a bounded sample of the workload (see bounded_oracle)."""
    from oracle import ref
    wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
    with oracle_prints_to_stderr():
        r, cores, t_pass, frame, shrink = bounded_oracle(ref, wl, pt, cfg)
        for _ in range(max(0, args.warmup - 2)):
            r.render(1)
        r.ray_count(reset=True)
        t0 = time.time()
        for _ in range(args.steps):
            r.render(1)
        dt = time.time() - t0
    rays = r.ray_count()
    value = rays / dt / 1e6
    sample = f"{args.steps} sample passes of the {frame} of {cfg['name']} (one step = 1 spp), OpenMP over rows on {cores} threads"
    return {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        # a step of the GPU arm is a frame of samples_per_frame passes of the whole image: the bounded sample's time, scaled to that size
        "ms_per_step": dt / args.steps * 1e3 * shrink * cfg["settings"].get("samples_per_frame", 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(cfg, args), "samples_per_px_per_s": args.steps / dt / shrink,
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def run_ours(args, cfg, pt):
    import ctypes as C
    import numpy as np
    import torch

    wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    W, H = cfg["width"], cfg["height"]
    mg = importlib.import_module("hiprt-path-tracer_b200.multigpu")
    restir = cfg["options"]["lss"] == "restir_di"
    sharding = args.sharding if args.sharding != "auto" else ("samples" if args.workload in ("cornell", "adaptive") else "columns")
    if restir and sharding != "samples":
        sharding = "tiles"              # reservoirs are per pixel and per frame: ReSTIR DI shards by row tiles + halo (or, on request, by sample passes: every rank its own reservoir history)
    tiles = mg.column_tiles(W, world) if sharding == "columns" else mg.row_tiles(H, world)
    halo = mg.restir_halo_rows() if restir else 0
    device = torch.device("cuda", local_rank)
    ctx = pt.Context(local_rank)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="gpu")
    lib = ctx.lib
    # The data-path collectives (frame gather, ReSTIR halo, accumulation all-reduce) are the C ABI's (include/b200pt_comm.h: NCCL bound
    # by libb200pt.so, enqueued on the render stream); torch.distributed is only the launcher's rendezvous for the 128-byte NCCL id
    comm = None
    if dist:
        ident = [mg.Comm.unique_id(lib) if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0, device=device)
        comm = mg.Comm(ctx, rank, world, ident[0])
        shard_kind = mg.Comm.COLUMNS if sharding == "columns" else mg.Comm.ROWS
        assert mg.Comm.tile(lib, shard_kind, W, H, world, rank) == ((tiles[rank][0], 0, tiles[rank][1], H) if sharding == "columns" else (0, tiles[rank][0], W, tiles[rank][1]))
    bvh_info = ctx.bvh_info(g.bvh)
    in_flight = int(os.environ.get("B200PT_SAMPLES_IN_FLIGHT", "16"))
    batched = g.renders_sample_batches

    def restart():
        # Sample-pass sharding: every rank draws its per-launch seeds from its own host stream, so the ranks render independent
        # samples at the SAME sample_number (the per-pixel streams are seeded by (pixel, sample_number, random_seed),
        # FullPathTracer.h:124-129; adaptive sampling's thresholds are sample numbers, so the sample number itself is not offset)
        if world > 1 and sharding == "samples":
            g.set_rng_seed(42 + 7919 * rank)
        g.reset()
        g.resize(W, H)
        wl.apply_settings(g.render_settings, cfg["settings"])
        if world > 1 and sharding == "tiles":
            # Tile sharding: rank r owns rows tiles[r]; global pixel indices and the shared seed sequence make its pixels
            # bit-identical to the 1-GPU render (tests/test_gpu_parity.py::test_tile_sharding_*, test_restir_di_tile_sharding_*)
            g.set_tile((0, tiles[rank][0], W, tiles[rank][1]), halo=halo)
        elif world > 1 and sharding == "columns":
            # the same with column bands: balanced on landscape frames (sky above, geometry below)
            g.set_tile((tiles[rank][0], 0, tiles[rank][1], H))

    def step():
        g.render_frame()
        if restir and world > 1:
            # ReSTIR DI halo: this frame's output reservoirs of the rows within 20 of a tile edge go to the neighbouring ranks
            # (36 B/px; the G-buffer of those rows is recomputed locally by CameraRays over tile + halo)
            comm.exchange_halo(g.device_pointer("restir_output_reservoirs"), H, W * 36, halo)     # on the render stream: ordered after the frame

    # Column bands of a landscape frame do not cost the same (sky columns end at the first miss, terrain columns bounce 8 times): before
    # anything is measured the ranks balance the plan on the device time of their own band, three frames, with the C ABI's
    # b200pt_comm_balanced_bounds (a tile only changes which rank renders a pixel, never its value: tests/test_gpu_parity.py::test_column_tile_*)
    balance_log = []
    if world > 1 and sharding == "columns" and not args.no_balance:
        bounds = [t[0] for t in tiles] + [W]
        evb0, evb1 = C.c_void_p(), C.c_void_p()
        lib.b200pt_event_create(ctx.handle, C.byref(evb0))
        lib.b200pt_event_create(ctx.handle, C.byref(evb1))
        for _ in range(3):
            restart()
            step()                     # allocations for the new band size
            g.synchronize()
            lib.b200pt_event_record(ctx.handle, evb0, None)
            step()
            lib.b200pt_event_record(ctx.handle, evb1, None)
            g.synchronize()
            ctx.sync()
            band_ms = C.c_float()
            lib.b200pt_event_elapsed_ms(ctx.handle, evb0, evb1, C.byref(band_ms))
            bounds, costs = comm.rebalance_columns(bounds, band_ms.value, W)
            balance_log.append({"ms": [round(c, 2) for c in costs], "bounds": list(bounds)})
            tiles = [(bounds[r], bounds[r + 1]) for r in range(world)]
        lib.b200pt_event_destroy(ctx.handle, evb0)
        lib.b200pt_event_destroy(ctx.handle, evb1)

    if args.profile:
        # under ncu: only what the timed pass launches (3 warm-up frames + K frames), so that the launch list is the step's
        restart()
        for _ in range(3 + args.steps):
            step()
        g.synchronize()
        print(json.dumps({"profile": cfg["name"], "frames": 3 + args.steps, "path_lanes": g.path_lanes}))
        return None

    # ---- 1. counting pass: the same frames as the timed pass, counting build ----
    # (ReSTIR DI passes counted as single kernels: their phased form queues a superset of the shadow rays -- the reference's ray count is the
    # metric's numerator -- and the queue-only counters then belong to FullPathTracer alone, as the trace stage time does)
    restart()
    ctx.counters_enable(True)
    ctx.set_restir_phases(False)
    for _ in range(args.warmup):
        step()
    g.synchronize()
    ctx.counters_read(True)
    for _ in range(args.steps):
        step()
    g.synchronize()
    total = ctx.counters_read(True)
    ctx.set_restir_phases(True)
    res = pt.structs.Int2(W, H)
    pt._capi.check(lib.b200pt_launch_CameraRays(ctx.handle, C.byref(g.render_data), res, None, None), "b200pt_launch_CameraRays")
    ctx.sync()
    cam = ctx.counters_read(True)
    ctx.counters_enable(False)
    rays_per_step = total["rays"] / args.steps
    fpt = {k: total[k] / args.steps - cam[k] for k in total}
    assert total["stack_overflows"] == 0, "traversal stack overflow"

    # ---- 1b. stage pass: the same frames with per-stage CUDA events inside FullPathTracer (timing build, untimed otherwise):
    #          gives the device time of the ray-queue traversal kernels, the kernels the node-fetch roofline is about ----
    restart()
    for _ in range(max(args.warmup, 3)):
        step()
    g.synchronize()
    lanes = g.path_lanes
    g.set_path_lanes(1)                  # the stages of concurrent lanes overlap: stage times are taken with a single lane
    ctx.path_profile_enable(True)
    for _ in range(args.steps):
        step()
    g.synchronize()
    stage_ms = {k: v / args.steps for k, v in ctx.path_profile_read(True).items()}
    ctx.path_profile_enable(False)
    g.set_path_lanes(lanes)

    # ---- 2. timed pass: timing build, CUDA events on the launching stream ----
    restart()
    for _ in range(max(args.warmup, 3)):
        step()
    g.synchronize()
    ev0, ev1 = C.c_void_p(), C.c_void_p()
    lib.b200pt_event_create(ctx.handle, C.byref(ev0))
    lib.b200pt_event_create(ctx.handle, C.byref(ev1))
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    kernel_ms = {"CameraRays": 0.0, "FullPathTracer": 0.0}
    lib.b200pt_event_record(ctx.handle, ev0, None)
    for _ in range(args.steps):
        step()
        g.synchronize()
        for k in kernel_ms:
            kernel_ms[k] += g.last_kernel_ms(k)
    lib.b200pt_event_record(ctx.handle, ev1, None)
    ctx.sync()
    ms = C.c_float()
    lib.b200pt_event_elapsed_ms(ctx.handle, ev0, ev1, C.byref(ms))
    elapsed_ms = ms.value
    reduce_ms = 0.0
    if dist:
        # once per OUTPUT, not per sample: the frame assembled on every rank (tile sharding) or the accumulation buffers summed into a
        # separate buffer (sample-pass sharding; multigpu.reduce_accumulation explains why not in place)
        d_pixels = g.device_pointer("pixels")
        d_sum = ctx.malloc(W * H * 12) if sharding == "samples" else None
        adaptive = sharding == "samples" and bool(cfg["settings"].get("enable_adaptive_sampling"))
        d_lum, d_cnt = (ctx.malloc(W * H * 4), ctx.malloc(W * H * 4)) if adaptive else (None, None)

        def collective():
            if sharding == "samples":
                comm.allreduce_sum(d_pixels, d_sum.value, W * H * 3, "f32")
                if adaptive:
                    comm.allreduce_sum(g.device_pointer("pixel_squared_luminance"), d_lum.value, W * H, "f32")
                    comm.allreduce_sum(g.device_pointer("pixel_sample_count"), d_cnt.value, W * H, "i32")
            else:
                comm.gather_tiles(d_pixels, W, H, 12, shard_kind)

        collective()                 # one untimed round first (NCCL channel set-up, staging allocation); the gathers are idempotent
        ctx.sync()
        comm.barrier()
        lib.b200pt_event_record(ctx.handle, ev0, None)
        collective()
        lib.b200pt_event_record(ctx.handle, ev1, None)
        ctx.sync()
        lib.b200pt_event_elapsed_ms(ctx.handle, ev0, ev1, C.byref(ms))
        reduce_ms = ms.value
        elapsed_ms += reduce_ms
        for p in (d_sum, d_lum, d_cnt):
            if p is not None:
                ctx.free(p)
        comm.barrier()
    torch.cuda.synchronize()
    clocks = sampler.summary() if sampler else None

    if dist:
        elapsed_ms = comm.allreduce_host([elapsed_ms], "max")[0]                      # device-timed, max over ranks
        total_rays = comm.allreduce_host([rays_per_step * args.steps], "sum")[0]
    else:
        total_rays = rays_per_step * args.steps

    # ---- 3. end to end through the host API with HOST buffers: camera in, accumulated frame out, every step ----
    restart()
    own_y0, own_rows = (tiles[rank][0], tiles[rank][1] - tiles[rank][0]) if (world > 1 and sharding == "tiles") else (0, H)
    pinned = C.c_void_p()
    lib.b200pt_host_alloc_pinned(ctx.handle, C.c_size_t(W * H * 12), C.byref(pinned))
    for _ in range(3):
        step()
    g.synchronize()
    lib.b200pt_event_record(ctx.handle, ev0, None)
    for _ in range(args.steps):
        g.set_camera(cfg["camera"])                       # host struct -> kernel arguments (H2D with each launch)
        step()
        lib.b200pt_memcpy_d2h(ctx.handle, pinned, C.c_void_p(g.device_pointer("pixels") + own_y0 * W * 12), C.c_size_t(W * own_rows * 12), None)
        g.synchronize()
    lib.b200pt_event_record(ctx.handle, ev1, None)
    ctx.sync()
    lib.b200pt_event_elapsed_ms(ctx.handle, ev0, ev1, C.byref(ms))
    e2e_ms = ms.value
    if dist:
        e2e_ms = comm.allreduce_host([e2e_ms], "max")[0]
    lib.b200pt_host_free_pinned(ctx.handle, pinned)

    if comm:
        comm.close()
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return None

    value = total_rays / (elapsed_ms * 1e-3) / 1e6
    peak, peak_kind = measured_hbm_peak()
    fpt_ms = kernel_ms["FullPathTracer"] / args.steps
    # algorithmic bytes per ray (SURVEY.md 8d / DESIGN.md): 32 B ray in + 16 B hit out + 80 B per node + 48 B per triangle,
    # over the rays / nodes / triangles of the ray-queue kernel ALONE (per-thread traversals inside other kernels are not credited to it)
    queue = {k: fpt["queue_" + k] for k in ("rays", "nodes", "triangles")}
    if queue["rays"] <= 0:
        queue = {k: fpt[k] for k in ("rays", "nodes", "triangles")}     # shapes without the ray-queue kernel: every traversal of FullPathTracer
    fpt_bytes = 48.0 * queue["rays"] + 80.0 * queue["nodes"] + 48.0 * queue["triangles"]
    trace_ms = stage_ms["trace"] if stage_ms["trace"] > 0 else fpt_ms    # bounce >= 1 rays + all NEE rays go through the ray-queue kernel
    achieved = fpt_bytes / (trace_ms * 1e-3) / 1e9
    # kernels per step (wavefront, phased MIS shape): CameraRays + per bounce {gen | trace + surface} + scan + scatter +
    # 2 per sampled light kind (eval, BSDF sample) + NEE any-hit trace + NEE closest trace (emissives only) + bsdf
    nb = cfg["settings"].get("nb_bounces", 8)
    has_lights, has_env = len(cfg["scene"].emissive) > 0, cfg["envmap"] is not None and cfg["options"]["ess"] != "none"
    # (shadow rays and NEE light rays of a bounce share one launch: k_trace_queue_pair)
    per_bounce = 2 + 2 * has_lights + 2 * has_env + (1 if (has_lights or has_env) else 0) + 1
    spf = cfg["settings"].get("samples_per_frame", 1)
    if batched:
        # fused sample batches: per lane and batch camera + trace + camera-surface, the bounce chain once for all passes in flight, accumulate
        batches = -(-spf // max(1, in_flight))
        launches_per_step = batches * lanes * (3 + per_bounce + nb * (2 + per_bounce) + 1)
    else:
        launches_per_step = spf * (1 + lanes * ((1 + per_bounce) + nb * (2 + per_bounce)) + (9 if restir else 0))      # every path lane launches its own band; phased ReSTIR DI: presampling + 5 (initial candidates: collect, trace, resolve, trace, visibility) + 3 (spatiotemporal: collect, trace, resolve) + CameraRays counted in the leading 1
    # DRAM bytes of the kernel per step and its time-weighted ncu counters (L2 throughput as a fraction of the L2 peak, hit rates, issue rate, lanes
    # per instruction), from the committed `ncu --set full` capture of the same command (profiles/traffic.json); null when not captured
    traffic, ncu_weighted = measured_traffic(args.workload, W, H, cfg["settings"].get("samples_per_frame", 1))
    out = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "strong" if (world > 1 and sharding in ("tiles", "columns")) else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {**workload_config(cfg, args), "parallelism": "single GPU" if world == 1 else ((f"row-tile sharding x{world}" + (f" + {halo}-row ReSTIR reservoir halo exchange" if restir else "") + ", frame gathered once per output") if sharding == "tiles"
                                                                      else (f"column-tile sharding x{world}" + (" (bands balanced on measured band times)" if balance_log else "") + ", frame gathered once per output; e2e reads the whole accumulation buffer back (a column band is strided)") if sharding == "columns"
                                                                      else f"sample-pass sharding x{world}, accumulation all-reduce") + ", scene+BVH replicated",
                   "l2": "inputs larger than L2: the per-step G-buffer round trip is 345 B/px = 715 MB at 1080p (L2 = 126 MB)",
                   "bvh": bvh_info, "path_lanes": lanes, "samples_in_flight": in_flight if batched else 0, "rays_per_step_per_gpu": rays_per_step, "allreduce_ms": reduce_ms,
                   **({"column_balance": {"band_ms_before": balance_log[0]["ms"], "band_ms_last_measured": balance_log[-1]["ms"], "bounds": balance_log[-1]["bounds"]}} if balance_log else {})},
        "samples_per_px_per_s": (1 if (world > 1 and sharding in ("tiles", "columns")) else world) * spf * args.steps / (elapsed_ms * 1e-3),
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
        "e2e": {"value": total_rays / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 2 * 1328 + 196, "d2h_bytes_per_step": W * own_rows * 12},
        "roofline": {"bound": "hbm", "frac_of": "HBM copy peak (the contract's denominator; the kernel is L2-latency-bound, not HBM-bound)", "kernel": "k_trace_queue (ray-queue BVH8 traversal inside FullPathTracer; per step = sum of its launches)", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "ncu": ncu_weighted, "l2_frac": (ncu_weighted["l2_throughput_pct"] / 100.0) if ncu_weighted else None,
                     "peak_source": peak_kind, "ms_per_launch": trace_ms, "full_path_tracer_ms": fpt_ms, "stage_ms": stage_ms,
                     "camera_rays_ms_per_launch": kernel_ms["CameraRays"] / args.steps,
                     "stage_ms_lanes": 1, "rays_per_launch": queue["rays"], "nodes_per_ray": queue["nodes"] / max(1.0, queue["rays"]),
                     "triangles_per_ray": queue["triangles"] / max(1.0, queue["rays"]), "algorithmic_bytes_per_step": fpt_bytes,
                     "limiter": "latency of dependent node / triangle fetches served by L1/L2 (ncu: long_scoreboard; DRAM traffic is several times below the algorithmic bytes), see DESIGN.md 4.2"},
    }

    # ---- 4. CPU baseline: the reference's own CPU kernels on a bounded sample of the same workload (rank 0, N = 1) ----
    if world == 1 and not args.no_cpu_baseline:
        from oracle import ref
        # Bounded sample (~10-30 s of CPU work), every host core (bounded_oracle)
        with oracle_prints_to_stderr():
            r, cores, t_pass, frame, shrink = bounded_oracle(ref, wl, pt, cfg)
            n = int(min(16, max(1, 12.0 / max(t_pass, 1e-3))))
            r.ray_count(reset=True)
            t0 = time.time()
            r.render(n)
            dt = time.time() - t0
        out["cpu_baseline"] = {"value": r.ray_count() / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                               "sample": f"{n} sample passes of the {frame} of {cfg['name']} with the reference's CPU kernels (oracle/_ref), OpenMP over rows",
                               "samples_per_px_per_s": n / dt / shrink}
    if dist:
        dist.destroy_process_group()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="synthetic", help="synthetic = C3 (10 M triangles + envmap, the config the target is quoted on; default), cornell = C2, many_lights = C4, adaptive = C5")
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--sharding", default="auto", choices=["auto", "samples", "tiles", "columns"],
                    help="N > 1: sample-pass sharding (weak scaling, default for cornell), column tiles (strong scaling, default otherwise) or row tiles (strong scaling; always for ReSTIR DI)")
    ap.add_argument("--samples-per-frame", type=int, default=16, help="sample passes per step (= per render() call): the reference's samples_per_frame render setting")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-balance", action="store_true", help="N > 1, column tiles: keep the uniform bands instead of balancing them on measured band times")
    ap.add_argument("--profile", action="store_true", help="for ncu: run only 3 warm-up + K frames of the timed pass (no counting / stage / e2e / CPU passes), print no bench line")
    args = ap.parse_args()

    pt = importlib.import_module("hiprt-path-tracer_b200")
    wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
    cfg = getattr(wl, args.workload)(args.width, args.height)
    cfg["settings"]["samples_per_frame"] = args.samples_per_frame

    if args.impl == "reference":
        if int(os.environ.get("RANK", 0)) != 0:
            return
        args.steps = min(args.steps, 8)
        print(json.dumps(run_reference(args, cfg, pt)))
        return
    out = run_ours(args, cfg, pt)
    if out is not None:
        print(json.dumps(out))


if __name__ == "__main__":
    main()
