"""bench.py — measures BASELINE.json's metric (Mpaths/s, with Mrays/s alongside) for the wavefront path-tracing core.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference ...                     (the CPU restatement of the reference, all host threads)

One "step" = one full render of the workload (default: C2 = bunny_5k conductor/PBR scene, 1920x1080 @ 256 spp,
max_depth 8, recurrence sampler, box r=0.5 — BASELINE.json configs[1]).  `value` is measured with the scene resident
in HBM and the framebuffer left on the device; `e2e` goes through spt_render with host buffers and re-sends the
flattened scene from pinned memory every step.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "simple-path-tracer_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs (B200_PROFILING.md recipe)."""

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def workload(args):
    from spt import scenes
    cfg = dict(scenes.CONFIGS[args.workload])
    w = args.width or cfg["width"]
    h = args.height or cfg["height"]
    spp = args.spp or cfg["spp"]
    return cfg, w, h, spp


def cpu_sample_size(w, h, cores):
    """Bounded sample of the workload for the CPU arm: the same scene, camera, aspect ratio, depth and sampler at reduced
    resolution, with the height a multiple of the reference's thread count.  PathTracer::render cuts the image into
    2 x cores contiguous row bands of height / threads rows and gives the LAST band the remainder (renderer/util.rs:6-19),
    and its wall time is that of the slowest band: a sample whose height is not a multiple of the thread count leaves a
    third of the rows on one thread (round 1 sampled 160 x 90 on 32 threads: 2 rows per band, 28 on the last) and
    measures that thread, not the machine.  At the real sizes the bands are 33 rows (last: 57) of 1080; here 4 rows each."""
    threads = 2 * cores
    rows = max(4, -(-32 // threads))  # 4 rows per band; more on a box with very few cores, so that the sample keeps a usable size
    sh = min(h, rows * threads)
    sh -= sh % threads if sh >= threads else 0
    sw = max(16, int(round(sh * w / h)))
    return sw, max(1, sh), threads


def cpu_run(osc, sw, sh, spp, seconds, seed=1):
    """Reference-faithful render (oracle rng_mode 0: per-thread streams and sampler state, 2 x cores row bands) of about `seconds`."""
    t = time.time(); _, st = osc.render(sw, sh, spp=1, rng_mode=0); probe = time.time() - t
    sspp = int(max(1, min(spp, seconds / max(probe, 1e-6))))
    t = time.time(); _, st = osc.render(sw, sh, spp=sspp, rng_mode=0, seed=seed); dt = time.time() - t
    return st, dt, sspp


def band_note(sh, threads):
    t = max(1, min(threads, sh))
    per = sh // t
    return f"{t} threads x {per} rows per band (last band {sh - per * (t - 1)} rows)"


def run_reference(args):
    """The reference arm: the reference's own CPU algorithm (oracle restatement — the Rust binary cannot be built here)
    on the box's host cores, reference-faithful mode (its BVH incl. Q1/Q2, its per-thread RNG/sampler discipline, 2 x cores
    threads on contiguous row bands)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import orc
    from spt import scenes
    cfg, w, h, spp = workload(args)
    fs = scenes.load(args.workload)
    osc = orc.OracleScene(fs)
    cores = os.cpu_count() or 1
    sw, sh, threads = cpu_sample_size(w, h, cores)
    st, probe_dt, sspp = cpu_run(osc, sw, sh, spp, args.ref_seconds)           # also the warm-up
    times, paths, rays = [], 0, 0
    for k in range(args.steps):
        t = time.time(); _, st = osc.render(sw, sh, spp=sspp, rng_mode=0, seed=k + 1); times.append(time.time() - t)
        paths += st["paths"]; rays += st["ext_rays"] + st["shadow_rays"] + st["probe_rays"]
    total = sum(times)
    v = paths / total / 1e6
    # the same algorithm on the checker's BVH (oracle fix_q1: the evidently intended bucket rule): how much of the distance to the
    # GPU number is the reference's degenerate tree (Q1: ~600x the triangle tests per ray on the bunny scenes)
    fst, fdt, fspp = cpu_run(orc.OracleScene(fs, fix_q1=True), sw, sh, spp, min(args.ref_seconds, 8.0))
    sample = (f"{args.workload} scene at {sw}x{sh} @ {sspp} spp per step (full config: {w}x{h} @ {spp} spp), reference BVH (Q1/Q2) + per-thread RNG, "
              + band_note(sh, threads))
    print(json.dumps({
        "impl": "reference", "metric": "Mpaths/s", "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {w}x{h} @ {spp} spp, max_depth 8", "sampler": "recurrence",
                   "same_config_note": "same scene / camera / aspect / depth / sampler as the GPU arm; resolution and spp reduced to a bounded sample (Mpaths/s is resolution- and spp-invariant up to band balance, which the sample keeps even)"},
        "mrays_per_s": rays / total / 1e6,
        "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": cores, "kind": "port", "sample": sample},
        "cpu_baseline_fixed_bvh": {"value": fst["paths"] / fdt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                                   "sample": f"same sample @ {fspp} spp on the checker BVH (oracle fix_q1: intended SAH bucket rule, identical hits), {fdt:.1f} s",
                                   "tris_per_ray": fst["tris_tested"] / max(1, fst["ext_rays"] + fst["shadow_rays"])},
        "tris_per_ray": st["tris_tested"] / max(1, st["ext_rays"] + st["shadow_rays"]),
        "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--ref-seconds", type=float, default=8.0, help="CPU seconds per reference/cpu_baseline step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip the secondary C5 @ 512 spp figure (about 10 s on one GPU)")
    ap.add_argument("--no-parity", action="store_true", help="skip the converged-image parity block (about 30 s of CPU oracle)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from spt import abi, core, scenes

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(local)
    W = max(3, args.warmup)
    cfg, w, h, spp = workload(args)
    fs = scenes.load(args.workload)
    ctx = core.Context(local)
    sc = core.Scene(ctx, fs)
    info = sc.info()
    dev = torch.device("cuda", local)
    fb = torch.zeros((h, w, 3), dtype=torch.float32, device=dev)
    kw = dict(spp=spp, max_depth=8, sampler="recurrence", rank=rank, world_size=world, pool_size=args.pool)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: the library's own NCCL communicator (spt_comm_create; torch.distributed only broadcasts the 128-byte id and provides the
    # barrier / timing reductions).  spt_render_gather renders this rank's tiles and moves only the owned tiles' integer film to rank 0.
    comm = core.Comm.from_torch(ctx) if world > 1 else None
    gather_ms = []

    def step(seed, flags=0):
        if world > 1:
            st = sc.render_gather(comm, fb.data_ptr(), w, h, root=0, seed=seed, flags=flags, **kw)
            gather_ms.append(1e3 * st["seconds_stage"][5])
        else:
            st = sc.render_device(fb.data_ptr(), w, h, seed=seed, flags=flags, **kw)
        return st

    for i in range(W):
        step(100 + i)
    # ---- timed region: device-resident inputs and outputs
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    barrier()
    gather_ms.clear()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record()  # spt_render_device returns after its own stream has drained, so events on this stream bracket the device work
    stats = [step(i + 1) for i in range(args.steps)]
    ev1.record()
    barrier()
    wall = time.perf_counter() - t0
    elapsed = ev0.elapsed_time(ev1) * 1e-3
    # per-rank breakdown of the timed region: device time inside spt_render_device (CUDA events on the library's stream)
    # and of the NCCL framebuffer reduce
    render_ms = 1e3 * sum(s_["seconds_total"] for s_ in stats) / args.steps
    reduce_ms = sum(gather_ms) / args.steps if gather_ms else 0.0
    brk = torch.tensor([render_ms, reduce_ms, 1e3 * wall / args.steps], dtype=torch.float64, device=dev)
    brk_all = [torch.zeros_like(brk) for _ in range(world)]
    if world > 1:
        dist.all_gather(brk_all, brk)
    else:
        brk_all = [brk]
    brk_list = [b.tolist() for b in brk_all]
    tt = torch.tensor([elapsed], dtype=torch.float64, device=dev)
    agg = torch.tensor([sum(s[k] for s in stats) for k in ("paths", "ext_rays", "shadow_rays", "probe_rays", "kernel_launches")],
                       dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
    elapsed = float(tt.item())
    paths, ext, shadow, probe, launches = (float(x) for x in agg.tolist())
    value = paths / elapsed / 1e6
    mrays = (ext + shadow + probe) / elapsed / 1e6

    # ---- e2e: host buffers in and out (scene re-sent from pinned memory, framebuffer read back) every step
    host = torch.empty((h, w, 3), dtype=torch.float32, pin_memory=True)  # page-locked result buffer, allocated once
    host_np = host.numpy()
    barrier()
    e2e_t0 = time.perf_counter()
    h2d = 0
    for i in range(args.steps):
        h2d = sc.upload()                      # scene arrays host (pinned staging) -> device
        if world > 1:
            step(i + 1)
            if rank == 0:
                host.copy_(fb)                 # reduced frame device -> host
        else:
            sc.render(w, h, seed=i + 1, out=host_np, **{k: v for k, v in kw.items()})  # spt_render copies the frame into `out`
    barrier()
    e2e_elapsed = time.perf_counter() - e2e_t0
    te = torch.tensor([e2e_elapsed], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = paths / float(te.item()) / 1e6
    if sampler:
        sampler.stop_flag = True
        sampler.join(timeout=2)

    # ---- secondary figure on the north star's own multi-GPU config (all ranks): C5 = 2000 instanced bunnies at 3840x2160, 512 of its
    # 4096 spp (paths/s is spp-invariant at this length: 4.2 G samples).  The headline stays C2 so that every N reports the same workload.
    scale_c5 = None
    if args.workload == "c2" and not args.no_secondary:
        c5cfg = scenes.CONFIGS["c5"]
        w5, h5, spp5 = c5cfg["width"], c5cfg["height"], 512
        sc5 = core.Scene(ctx, scenes.load("c5"))
        fb5 = torch.zeros((h5, w5, 3), dtype=torch.float32, device=dev)
        kw5 = dict(spp=spp5, max_depth=8, sampler="recurrence")

        def step5(seed):
            if world > 1:
                return sc5.render_gather(comm, fb5.data_ptr(), w5, h5, root=0, seed=seed, **kw5)
            return sc5.render_device(fb5.data_ptr(), w5, h5, seed=seed, **kw5)

        step5(50)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st5 = step5(51)
        e1.record()
        barrier()
        t5 = torch.tensor([e0.elapsed_time(e1) * 1e-3], dtype=torch.float64, device=dev)
        a5 = torch.tensor([st5["paths"], st5["ext_rays"] + st5["shadow_rays"] + st5["probe_rays"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t5, op=dist.ReduceOp.MAX)
            dist.all_reduce(a5, op=dist.ReduceOp.SUM)
        scale_c5 = {"workload": f"c5: {w5}x{h5} @ {spp5} of its 4096 spp, 2000 instances x bunny_5k = 10.0 M instanced triangles, max_depth 8",
                    "value": float(a5[0].item()) / float(t5.item()) / 1e6, "unit": "Mpaths/s", "mrays_per_s": float(a5[1].item()) / float(t5.item()) / 1e6,
                    "ms_per_frame": 1e3 * float(t5.item()), "steps": 1, "warmup": 1, "n_gpus": world}
        sc5.close()
        del fb5

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (extend = closest-hit traversal), rank 0
    # One extra step with per-stage CUDA events on the library's stream (it synchronises every iteration, so it is
    # kept out of the timed region), and one short counting step for the mean nodes / triangles / instances per ray.
    peaks, peak_src = load_peaks()
    s0 = sc.render_device(fb.data_ptr(), w, h, seed=1, flags=abi.RENDER_TIME_STAGES, **kw)
    cst = sc.render_device(fb.data_ptr(), w, h, seed=1, flags=abi.RENDER_COUNT_TRAVERSAL,
                           **dict(kw, spp=max(1, min(spp, 8))))
    n_rays = cst["ext_rays"] + cst["shadow_rays"] + cst["probe_rays"]
    n_node, n_tri, n_inst = cst["nodes_visited"] / n_rays, cst["tris_tested"] / n_rays, cst["insts_visited"] / n_rays
    b_ray = 52.0 + 80.0 * n_node + 48.0 * n_tri + 64.0 * n_inst          # SURVEY §8d algorithmic bytes per traced ray
    ext_launches = s0["iterations"]
    ext_seconds = s0["seconds_stage"][1]
    rays_per_launch = s0["ext_rays"] / max(1, ext_launches)
    achieved = rays_per_launch * b_ray / max(ext_seconds / max(1, ext_launches), 1e-12) / 1e9
    # DRAM traffic and issue statistics of k_extend cannot be measured from inside a run (they need the profiler): they come from the
    # committed `ncu --set full` capture of one steady-state launch of this same command (profiles/r02_ncu_extend.json, made by
    # tools/gpu_profile.sh + tools/ncu_extract.py; it names the commit it was taken at).  `traffic` = its DRAM bytes per ray x the rays
    # per launch of THIS run; everything else in the roofline block is measured live.
    traffic, cap = None, None
    for name in ("r02_ncu_extend.json", "r01_ncu_extend_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as fh:
                cap = json.load(fh)
            traffic = cap["dram_bytes_per_launch"] / cap["rays_in_captured_launch"] * rays_per_launch
            cap["file"] = "profiles/" + name
            break
        except (OSError, KeyError):
            cap = None
    # regime (i) of SURVEY §8d: the scene is L2-resident, so the node / triangle / instance fetches are bounded by the
    # L2 -> SM read bandwidth, measured here with a 16-byte-load microbenchmark over a 32 MB read-only set
    l2 = None
    try:
        import ctypes
        bl = ctypes.CDLL(os.path.join(ROOT, "simple-path-tracer_b200", "lib", "libspt_benchkernels.so"))
        bl.spt_bench_l2_read_gbs.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        g = ctypes.c_double()
        if bl.spt_bench_l2_read_gbs(local, 32 << 20, 64, 5, ctypes.byref(g)) == 0:
            scene_bytes_per_ray = 80.0 * n_node + 48.0 * n_tri + 64.0 * n_inst
            l2_achieved = rays_per_launch * scene_bytes_per_ray / max(ext_seconds / max(1, ext_launches), 1e-12) / 1e9
            l2 = {"peak_gbs": g.value, "peak_source": "ld.global.cg.v4 over a 32 MB set, best of 5 (simple-path-tracer_b200/bench_kernels/l2_peak.cu)",
                  "achieved_gbs": l2_achieved, "frac": l2_achieved / g.value,
                  "what": "algorithmic node + triangle + instance bytes per ray x rays/s of k_extend (requests that hit L1 never reach L2)"}
    except OSError:
        pass
    # the library's default (spt_api.cu): samples/6 per rank within [2^22, 2^25] at first, widened after three iterations to
    # what finishes the render in ~10 iterations, at most samples/4 (allocated)
    pool_lo = min(1 << 22, max(1 << 16, w * h * spp))
    pool_slots = args.pool or min(1 << 25, max(pool_lo, w * h * spp // world // 6))
    pool_alloc = args.pool or min(1 << 25, max(pool_lo, w * h * spp // world // 4))
    roofline = {"bound": "hbm", "kernel": "k_extend", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / peaks["hbm_gbs"], "traffic": traffic, "peak_source": peak_src,
                "bytes_per_ray": b_ray, "nodes_per_ray": n_node, "tris_per_ray": n_tri, "insts_per_ray": n_inst,
                "rays_per_launch": rays_per_launch, "avg_launch_ms": 1e3 * ext_seconds / max(1, ext_launches),
                "grays_per_s": rays_per_launch / max(ext_seconds / max(1, ext_launches), 1e-12) / 1e9,
                "stage_seconds": {"extend": s0["seconds_stage"][1], "shade": s0["seconds_stage"][2], "connect": s0["seconds_stage"][3],
                                  "total": s0["seconds_total"]},
                "traffic_source": None if cap is None else {"file": cap["file"], "commit": cap.get("commit"), "dram_bytes_per_ray": cap["dram_bytes_per_launch"] / cap["rays_in_captured_launch"],
                                                            "how": "ncu --set full capture of one steady-state k_extend launch, scaled per ray to this run's rays per launch (not measured in this run)"},
                "issue": None if cap is None or "lanes_per_inst" not in cap else {
                    "what": "the binding roof: instruction issue (same ncu capture)", "commit": cap.get("commit"), "issue_active_pct": cap.get("issue_active_pct"),
                    "lanes_per_inst": cap.get("lanes_per_inst"), "alu_pipe_pct": cap.get("alu_pipe_pct"), "xu_pipe_pct": cap.get("xu_pipe_pct"),
                    "fma_pipe_pct": cap.get("fma_pipe_pct"), "warps_active_pct": cap.get("warps_active_pct"), "l1_hit_pct": cap.get("l1_hit_pct")},
                "l2": l2,
                "note": "scene working set (%.2f MB) is L2-resident: the HBM figure is the contract's denominator; neither memory roof binds, the kernel is bound by instruction issue (see `issue`)" % (info["bytes_total"] / 1e6)}
    out = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": W,
        "ms_per_step": 1e3 * elapsed / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {w}x{h} @ {spp} spp, max_depth 8, recurrence sampler, box r=0.5",
                   "scene": "bunny_5k x3 (conductor, pbr_metallic iso+aniso) + room + emissive quad, power_is MIS" if args.workload == "c2" else args.workload,
                   "parallelism": f"tiles{world}", "pool_slots": pool_slots, "pool_slots_allocated": pool_alloc,
                   "l2_policy": "per-step working set (path pool %d MB + film %d MB) exceeds the 126 MB L2; no explicit flush" % (pool_alloc * 184 >> 20, w * h * 24 >> 20)},
        "mrays_per_s": mrays, "rays_per_path": (ext + shadow + probe) / paths,
        "shadow_rays_culled_per_path": sum(s["shadow_culled"] for s in stats) / max(1.0, sum(s["paths"] for s in stats)),
        "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(w * h * 3 * 4)},
        "gpu_launches": int(launches), "scale_c5": scale_c5, "roofline": roofline, "clocks": sampler.summary() if sampler else None,
        "timing": {"source": "CUDA events bracketing the K steps, max over ranks", "wall_ms_per_step": float(max(b[2] for b in brk_list)),
                   "render_device_ms_per_rank": [round(float(b[0]), 3) for b in brk_list],
                   "gather_ms_per_rank": [round(float(b[1]), 3) for b in brk_list],
                   "gather": "spt_render_gather: owned tiles only, int64 film, grouped ncclSend/ncclRecv to rank 0 + unpack + resolve (device time per rank)"},
    }
    if world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import orc
        cores = os.cpu_count() or 1
        sw, sh, threads = cpu_sample_size(w, h, cores)
        st, dt, sspp = cpu_run(orc.OracleScene(fs), sw, sh, spp, 15.0)
        out["cpu_baseline"] = {"value": st["paths"] / dt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                               "sample": f"{args.workload} scene at {sw}x{sh} @ {sspp} spp ({st['paths']} paths, {dt:.1f} s), reference BVH (Q1/Q2) + per-thread RNG, " + band_note(sh, threads),
                               "mrays_per_s": (st["ext_rays"] + st["shadow_rays"] + st["probe_rays"]) / dt / 1e6,
                               "tris_per_ray": st["tris_tested"] / max(1, st["ext_rays"] + st["shadow_rays"])}
        fst, fdt, fspp = cpu_run(orc.OracleScene(fs, fix_q1=True), sw, sh, spp, 8.0)
        out["cpu_baseline_fixed_bvh"] = {"value": fst["paths"] / fdt / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": "port",
                                         "sample": f"same sample @ {fspp} spp ({fdt:.1f} s) on the checker BVH (oracle fix_q1: the intended SAH bucket rule, identical hits): the algorithm-equal CPU line",
                                         "tris_per_ray": fst["tris_tested"] / max(1, fst["ext_rays"] + fst["shadow_rays"])}
    if world == 1 and not args.no_parity:
        # converged-image parity of this workload against the reference-faithful oracle (the checker; outside every timed region)
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import parity_stats
        if args.workload in parity_stats.CONVERGED:
            out["parity"] = parity_stats.converged_parity(sc, fs, args.workload, check_one_call=False)
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
