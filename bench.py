#!/usr/bin/env python
"""Headline benchmark: rigid body-steps/s of the CUDA step, with the reference's own CPU implementation timed beside it.

  python bench.py --gpus N --steps K --warmup W            -> our arm, one JSON line (rank 0)
  python bench.py --impl reference --steps K --warmup W    -> the reference (oracle/_ref/oracle_step) on the host cores

Workload (every N, including 1): C5 of BASELINE.json -- 4 096 independent PhysicsServer3D spaces x 256 mixed bodies
(1 048 576 bodies), sharded by space over the ranks with no data-path collective ("scaling": "strong").  It is the config
inside the metric's 100 k - 1 M range that shards and the one the 1 M-bodies-at-60 Hz target is stated on.
At N = 1 the line also carries the settled C2 (100 k spheres) and C3 (1 M-body single pile) device numbers under
`config.also`.

A "step" is one PhysicsServer3D::step(1/60) of every space.  Regime: the scene is first SETTLED by an untimed prelude of a
fixed number of steps (independent of --warmup), so the timed steps are the steady state (bodies piled up, contact caches
warm), not the cheap free fall.  The reference arm starts from the same settled state: the prelude's body state is handed to
oracle_step (--state), which then takes its own --warmup steps to build its contact caches before its timed steps.
`value` is measured with the state resident in HBM (CUDA events around each step, L2 flushed between steps); `e2e` goes
through the C-ABI with host buffers: every step uploads the per-body applied forces from pinned host memory and reads every
body transform back (what the engine's state sync does).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from bradot_b200 import scenes  # noqa: E402

ORACLE = os.path.join(ROOT, "oracle", "_ref", "oracle_step")
METRIC = "rigid_body_steps_per_s"
UNIT = "body-steps/s"
# untimed settle prelude per scene (steps): long enough for the bodies to have landed and piled up
SETTLE = {"c1": 120, "c2": 150, "c3": 400, "c4": 120, "c5": 120}
REFERENCE_SAMPLE_SPACES = 256   # spaces of C5 the CPU arm steps (it steps spaces one after the other: same rate per body)


def build_scene(name, scale, rank=0, world=1):
    if name == "c5":
        total = max(world, int(round(4096 * scale)))
        lo, hi = shard_range(total, rank, world)
        return scenes.batched_worlds(hi - lo, seed0=1000 + lo), "C5: %d independent spaces x 256 mixed bodies, sharded by space" % total
    if name == "c2":
        if scale >= 1.0:
            return scenes.spheres_container(50, 40, 50), "C2: 100k spheres r=0.25 on a 50x40x50 lattice falling into 5 static planes"
        n = max(2, int(round(50 * scale ** (1 / 3.0))))
        return scenes.spheres_container(n, max(2, int(n * 0.8)), n, half=max(2.0, n * 0.3)), "C2 scaled"
    if name == "c1":
        return scenes.pyramid(1000), "C1: 1000-box pyramid on a static plane"
    if name == "c3":
        n = max(3, int(round(100 * scale ** (1 / 3.0))))
        return scenes.mixed_pile(n), "C3: %d mixed box/sphere/capsule bodies" % n ** 3
    if name == "c4":
        chains = max(1, int(round(25000 * scale)))
        return scenes.convex_chains(chains), "C4: %d convex hulls (64 distinct) in %d chains of 8 linked alternately by pin and hinge joints" % (chains * 8, chains)
    raise SystemExit("unknown scene " + name)


def shard_range(n_units, rank, world):
    """Contiguous block of independent units (spaces) owned by `rank`; sizes differ by at most one."""
    base, rem = divmod(n_units, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_max_ms(ms, device="cuda"):
    """max over ranks of a duration (the job is as slow as its slowest rank)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_process_info(stats, device="cuda", stream=None):
    """get_process_info of the whole job: all-reduce(SUM) of (active_objects, collision_pairs, island_count) over the ranks
    (reference: BradotPhysicsServer3D::step sums them over its spaces, bradot_physics_server_3d.cpp:1632-1648).  The one
    collective of the path besides the timing reduction; on CUDA it is issued on a side stream, off the step's stream."""
    import torch
    import torch.distributed as dist
    vals = [stats["active_objects"], stats["collision_pairs"], stats["island_count"]]
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return dict(zip(("active_objects", "collision_pairs", "island_count"), (int(v) for v in vals)))
    if stream is not None:
        with torch.cuda.stream(stream):
            t = torch.tensor(vals, dtype=torch.int64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        stream.synchronize()
    else:
        t = torch.tensor(vals, dtype=torch.int64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    a, c, i = (int(v) for v in t.tolist())
    return {"active_objects": a, "collision_pairs": c, "island_count": i}


def reduce_sum(x, device="cuda"):
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def n_dynamic(sc):
    return int(np.sum(sc.bodies["mode"] >= scenes.MODE_RIGID))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.rows, self.stop_flag, self.index = [], False, index

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                r = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                if r.returncode == 0 and r.stdout.strip():
                    self.rows.append([x.strip() for x in r.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(self.rows)}


def settle_on_gpu(sc, steps, device=0, stream=None):
    """The untimed prelude: `steps` steps of the scene on the GPU; returns the server (state resident) and the settled body
    state (n_bodies, 18) = transform, linear velocity, angular velocity, for the reference arm."""
    from bradot_b200 import server
    ps = server.PhysicsServerCUDA(device, stream)
    ps.load_scene(sc)
    for _ in range(steps):
        ps.step(sc.dt)
    ps.sync()
    state = ps.bodies_read_state()[:, 0:18].copy()
    return ps, state


def reference_sample(name, sc, state):
    """The bounded sample of the workload the CPU arm steps: for C5 its first REFERENCE_SAMPLE_SPACES spaces (the reference
    steps spaces one after the other, bradot_physics_server_3d.cpp:1642-1647, so its rate per body does not depend on how
    many there are); the whole scene otherwise."""
    if name != "c5" or sc.n_spaces <= REFERENCE_SAMPLE_SPACES:
        return sc, state, "the whole scene"
    keep = sc.bodies["space"] < REFERENCE_SAMPLE_SPACES
    sub = scenes.Scene(REFERENCE_SAMPLE_SPACES, sc.dt, sc.iterations)
    sub.shapes, sub.verts = sc.shapes, sc.verts      # shape indices stay valid
    sub.bodies = sc.bodies[keep].copy()
    return sub, (state[keep] if state is not None else None), "the first %d of its spaces (%d bodies)" % (REFERENCE_SAMPLE_SPACES, int(keep.sum()))


def run_reference(sc, steps, warmup, state=None, threads=-1):
    """Time the reference's own step on the host cores; returns the parsed JSON of oracle_step."""
    if not os.path.exists(ORACLE):
        raise SystemExit("oracle/_ref/oracle_step is missing (python oracle/build_ref.py builds it where /root/reference exists)")
    d = tempfile.mkdtemp(prefix="bpc_bench_")
    path = os.path.join(d, "scene.bscn")
    sc.write(path)
    cmd = [ORACLE, path, "--steps", str(steps), "--warmup", str(warmup), "--threads", str(threads)]
    spath = None
    if state is not None:
        spath = os.path.join(d, "state.bin")
        with open(spath, "wb") as f:
            f.write(np.uint32(len(state)).tobytes())
            f.write(np.ascontiguousarray(state, np.float32).tobytes())
        cmd += ["--state", spath]
    r = subprocess.run(cmd, capture_output=True, text=True)
    os.remove(path)
    if spath:
        os.remove(spath)
    os.rmdir(d)
    if r.returncode != 0:
        raise SystemExit("oracle_step failed: " + r.stderr[-1000:])
    return json.loads(r.stdout.strip().splitlines()[-1])


def cpu_leg(sub, sub_state, what, steps, warmup, settle):
    """(value, oracle info, cpu_baseline dict) of the reference on `sub`, a bounded sample of the settled scene."""
    info = run_reference(sub, steps, warmup, sub_state)
    nd = n_dynamic(sub)
    value = nd * info["steps"] / info["wall_s"]
    regime = "settled by %d untimed GPU steps, state handed over (--state)" % settle if sub_state is not None else "from the scene's initial state"
    return value, info, {
        "value": value, "unit": UNIT, "cores": info["threads"], "kind": "reference",
        "sample": "%s, %s; %d warm-up + %d timed steps, WorkerThreadPool with all host threads (oracle/_ref/oracle_step = the "
                  "reference's own code)" % (what, regime, warmup, info["steps"]),
        "ms_per_step": 1e3 * info["wall_s"] / info["steps"], "phase_us": info["phase_us"]}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, label = build_scene(args.scene, args.scale)
    settle = SETTLE[args.scene] if args.settle is None else args.settle
    sub, _, what = reference_sample(args.scene, sc, None)
    state = None
    try:    # the settled start state comes from the same untimed prelude as the GPU arm's (not part of the timed path)
        import torch
        if settle > 0 and torch.cuda.is_available():
            ps, state = settle_on_gpu(sub, settle)
            ps.finish()
    except Exception as e:  # no usable GPU for the prelude: time the reference from the scene's initial state
        sys.stderr.write("settle prelude unavailable (%s): the reference starts from the initial state\n" % e)
        state = None
    value, info, base = cpu_leg(sub, state, what, args.steps, args.warmup, settle)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * info["wall_s"] / info["steps"], "higher_is_better": True,
        "scaling": "strong" if args.scene == "c5" else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": label, "bodies": n_dynamic(sc), "dt": float(sc.dt), "solver_iterations": sc.iterations,
                   "regime": "settled: %d untimed steps before warm-up" % settle},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def solve_bytes_per_step(sc, pc, cbar, iters):
    """pre-solve 296 + 76 c bytes per pair, each solver iteration 296 + 108 c bytes per pair; a
    joint-iteration 296 B + 64 B per Jacobian row (pin 3, hinge 6)."""
    b = pc * ((296 + 76 * cbar) + iters * (296 + 108 * cbar))
    if len(sc.joints):
        rows = np.where(sc.joints["type"] == 0, 3, 6).astype(np.float64)
        b += float(np.sum(iters * (296 + 64 * rows)))
    return b


def measure_device(ps, sc, steps, flush, barrier):
    """Time `steps` steps of a settled server with CUDA events (L2 flushed before each); returns per-step averages."""
    import torch
    ps.set_profiling(True)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    keys = ("contact_pairs", "active_contacts", "collision_pairs", "colors", "full_pairs", "barrier_constraints", "kernel_launches")
    acc = dict.fromkeys(keys, 0)
    first = last = None
    barrier()
    for i in range(steps):
        flush.zero_()
        ev[i][0].record()
        ps.step(sc.dt)
        ev[i][1].record()
        last = ps.stats()    # synchronises; outside the event pair
        first = first or last
        for k in keys:
            acc[k] += last[k]
    barrier()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    stage_ms, stage_steps = ps.stage_times()
    ps.set_profiling(False)
    k = max(1, steps)
    return {"ms": ms, "avg": {q: acc[q] / k for q in keys}, "first": first, "last": last,
            "stage_ms": {s: v / max(1, stage_steps) for s, v in stage_ms.items()}}


def stage_report(sc, nb, m, iters, peak):
    """roofline of the solve stage + achieved GB/s of broadphase / narrowphase from the algorithmic bytes of SURVEY.md 8(d)."""
    avg, st = m["avg"], m["stage_ms"]
    pc = avg["contact_pairs"]
    cbar = avg["active_contacts"] / max(1.0, pc)
    sb = solve_bytes_per_step(sc, pc, cbar, iters)
    solve_ms = st["solve"]
    achieved = sb / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else 0.0
    # broadphase: 260 B per body + 16 B per pair; narrowphase: ~250 B per pair the decision pass tests (2 transforms + 2 shapes
    # + header) + 664 B per pair that reaches the full pass (manifold read + written)
    bp_bytes = 260.0 * nb + 16.0 * avg["collision_pairs"]
    np_bytes = 250.0 * avg["collision_pairs"] + 664.0 * avg["full_pairs"]
    bp_ms, np_ms = st["broadphase"], st["narrowphase"]
    gbs = {"broadphase": bp_bytes / (bp_ms * 1e-3) / 1e9 if bp_ms > 0 else None,
           "narrowphase": np_bytes / (np_ms * 1e-3) / 1e9 if np_ms > 0 else None,
           "broadphase+narrowphase": (bp_bytes + np_bytes) / ((bp_ms + np_ms) * 1e-3) / 1e9 if bp_ms + np_ms > 0 else None,
           "note": "algorithmic bytes (SURVEY.md 8d: 260 B/body + 16 B/pair; 250 B per pair tested + 664 B per pair through the "
                   "full pass) / stage time (CUDA events), this rank"}
    total = sum(st.values())
    roof = {"bound": "hbm", "kernel": "k_presolve + k_sweeps / k_solve_bins (pre-solve + %d Gauss-Seidel sweeps)" % iters,
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "avg_launch_ms": solve_ms,
            "algorithmic_bytes_per_launch": sb, "share_of_step": solve_ms / total if total > 0 else None}
    return roof, gbs, cbar


def traffic_for(scene):
    """DRAM bytes of one solve launch from the committed ncu capture of this scene's timed regime (profiles/)."""
    try:
        with open(os.path.join(ROOT, "profiles", "k_solve_traffic.json")) as f:
            t = json.load(f).get(scene)
        if t:
            return t["traffic_bytes"], "ncu --set full capture at step %d (%d colliding pairs): %s" % (t["step"], t["contact_pairs"], t["capture"])
    except (OSError, ValueError, KeyError):
        pass
    return None, None


def also_record(name, device, stream, flush, steps, peak):
    """Device numbers of another BASELINE config in its settled regime (N = 1 only)."""
    import torch
    sc, label = build_scene(name, 1.0)
    ps, _ = settle_on_gpu(sc, SETTLE[name], device, stream)
    for _ in range(3):
        ps.step(sc.dt)
    ps.sync()
    m = measure_device(ps, sc, steps, flush, torch.cuda.synchronize)
    nb, nd = ps.body_count(), n_dynamic(sc)
    ps.finish()
    roof, gbs, cbar = stage_report(sc, nb, m, sc.iterations, peak)
    ms = m["ms"] / steps
    return {"workload": label, "bodies": nd, "regime": "settled: %d untimed steps, then 3 warm-up + %d timed" % (SETTLE[name], steps),
            "ms_per_step": ms, "body_steps_per_s": nd / (ms * 1e-3), "hz": 1e3 / ms, "stage_ms_per_step": m["stage_ms"],
            "avg_broadphase_pairs": m["avg"]["collision_pairs"], "avg_contact_pairs": m["avg"]["contact_pairs"],
            "avg_colors": m["avg"]["colors"], "solve_roofline_frac": roof["frac"], "solve_achieved_gbs": roof["achieved"]}


def cuda_arm(args):
    import torch
    import torch.distributed as dist
    from bradot_b200 import server

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "INFO")          # the communicator (ranks, NVLS / ring) shows in the log
        os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc, label = build_scene(args.scene, args.scale, rank, world)
    settle = SETTLE[args.scene] if args.settle is None else args.settle
    nd = n_dynamic(sc)
    total_nd = reduce_sum(float(nd))
    stream = torch.cuda.current_stream().cuda_stream
    side = torch.cuda.Stream()                                 # the stats collective runs here, off the step's stream
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")  # 256 MiB > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident arm: settle (untimed, fixed), warm up, time
    ps, settled = settle_on_gpu(sc, settle, local, stream)
    for _ in range(args.warmup):
        ps.step(sc.dt)
    ps.sync()
    sampler = ClockSampler(local)
    sampler.start()
    m = measure_device(ps, sc, args.steps, flush, barrier)
    process_info = reduce_process_info(m["last"], stream=side)
    sampler.stop_flag = True
    sampler.join()
    ms_max = reduce_max_ms(m["ms"])
    value = total_nd * args.steps / (ms_max * 1e-3)
    nb = ps.body_count()
    ps.finish()

    # ---------------- end-to-end arm: host buffers through the C ABI, fresh context, same schedule
    ps2, _ = settle_on_gpu(sc, settle, local, stream)
    forces_t = torch.zeros((nb, 6), dtype=torch.float32).pin_memory()   # pinned host buffers, as an engine would keep
    xf_t = [torch.zeros((nb, 12), dtype=torch.float32).pin_memory() for _ in range(2)]
    forces, xf_host = forces_t.numpy(), [x.numpy() for x in xf_t]
    for _ in range(args.warmup):
        ps2.step(sc.dt)
    ps2.sync()
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    checksum = 0.0
    # every step: H2D of the force rows, the step, D2H of every transform.  The read-back is the pipelined one of the C ABI
    # (bpc_bodies_read_begin / _end, two pinned buffers): step k's transforms cross the bus while step k + 1 runs and are
    # consumed one step later -- what an engine's flush_queries does with last tick's state.
    for k in range(args.steps):
        ps2.bodies_write(server.FIELD_APPLIED_FORCE, forces)     # H2D: per-body force/torque accumulators
        ps2.step(sc.dt)
        ps2.bodies_read_end()                                      # step k - 1's transforms are on the host now
        if k:
            checksum += float(xf_host[(k - 1) & 1][-1, 10])
        ps2.bodies_read_begin(server.FIELD_TRANSFORM, xf_host[k & 1])   # D2H: every body transform (state sync)
    ps2.bodies_read_end()
    checksum += float(xf_host[(args.steps - 1) & 1][-1, 10])
    e1.record()
    barrier()
    e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    e2e_value = total_nd * args.steps / (reduce_max_ms(e2e_ms) * 1e-3)
    ps2.finish()

    if rank == 0:
        k = max(1, args.steps)
        iters = sc.iterations
        peak, peak_src = peaks()
        roof, gbs, cbar = stage_report(sc, nb, m, iters, peak)
        traffic, traffic_src = traffic_for(args.scene if args.scale >= 1.0 and world == 1 else "")
        roof.update({"traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src})
        avg = m["avg"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / k, "higher_is_better": True, "scaling": "strong" if args.scene == "c5" else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": label, "bodies": int(total_nd), "dt": float(sc.dt), "solver_iterations": iters,
                       "regime": "settled: %d untimed steps before warm-up; %d colliding pairs of %d broadphase pairs on this rank at the first "
                                 "timed step" % (settle, m["first"]["contact_pairs"], m["first"]["collision_pairs"]),
                       "parallelism": ("spaces sharded over %d ranks, no data-path collective; get_process_info by one all-reduce on a side "
                                       "stream" % world if args.scene == "c5" else "replicas only (%d independent copies)" % world) if world > 1 else "1 GPU",
                       "l2": "256 MiB buffer written between timed steps (L2 flush)",
                       "avg_broadphase_pairs": avg["collision_pairs"], "avg_contact_pairs": avg["contact_pairs"], "avg_contacts_per_pair": cbar,
                       "avg_full_narrowphase_pairs": avg["full_pairs"], "avg_colors": avg["colors"],
                       "avg_barrier_path_constraints": avg["barrier_constraints"], "hz": 1e3 * k / ms_max},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(forces.nbytes), "d2h_bytes_per_step": int(nb * 12 * 4),
                    "checksum": checksum},
            "gpu_launches": int(avg["kernel_launches"] * k),
            "process_info": process_info,
            "clocks": sampler.summary(),
            "roofline": roof,
            "stage_ms_per_step": m["stage_ms"],
            "stage_gbs": gbs,
        }
        if world == 1 and args.also and args.scene == "c5" and args.scale >= 1.0:
            line["config"]["also"] = {}
            for name in ("c2", "c3"):
                try:
                    line["config"]["also"][name] = also_record(name, local, stream, flush, args.also_steps, peak)
                except Exception as e:   # a failed side measurement must not lose the headline line
                    line["config"]["also"][name] = {"error": str(e)[:200]}
        if args.cpu_baseline and world == 1:
            kc = max(1, min(args.steps, args.cpu_steps))
            sub, sub_state, what = reference_sample(args.scene, sc, settled)
            _, _, line["cpu_baseline"] = cpu_leg(sub, sub_state, what, kc, min(args.warmup, 5), settle)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--scene", default="c5", choices=sorted(SETTLE))
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--settle", type=int, default=None, help="untimed settle steps (default: per scene)")
    ap.add_argument("--cpu-steps", type=int, default=10)
    ap.add_argument("--also-steps", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    ap.add_argument("--no-also", dest="also", action="store_false")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        cuda_arm(args)


if __name__ == "__main__":
    main()
