#!/usr/bin/env python
"""Headline benchmark: rigid body-steps/s of the CUDA step on the C2 scene (BASELINE.json configs[1]:
100k spheres falling into a static box container), with the reference's own CPU implementation timed
beside it.

  python bench.py --gpus 1 --steps K --warmup W             -> our arm, one JSON line
  python bench.py --impl reference --steps K --warmup W     -> the reference (oracle/_ref/oracle_step) on the host cores

A "step" is one PhysicsServer3D::step(1/60) of the whole scene.  `value` is measured with the state
resident in HBM (CUDA events around each step, L2 flushed between steps); `e2e` goes through the
C-ABI with host buffers: every step uploads the per-body applied forces from host memory and reads
all body transforms back (what the engine's state sync does).  The single-space scenes do not shard,
so N > 1 runs N independent replicas, one per rank ("replicas only"); the only collective
is the max-over-ranks of the timing.
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


def reduce_process_info(stats, device="cuda"):
    """get_process_info of the whole job: all-reduce(SUM) of (active_objects, collision_pairs, island_count) over the ranks
    (reference: BradotPhysicsServer3D::step sums them over its spaces, bradot_physics_server_3d.cpp:1632-1648).  The only
    collective besides the timing reduction, issued after the timed region."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([stats["active_objects"], stats["collision_pairs"], stats["island_count"]], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
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
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(self.rows)}


def run_reference(args, sc, label, steps, warmup, threads=-1):
    """Time the reference's own step on the host cores; returns the parsed JSON of oracle_step."""
    if not os.path.exists(ORACLE):
        raise SystemExit("oracle/_ref/oracle_step is missing (python oracle/build_ref.py builds it where /root/reference exists)")
    d = tempfile.mkdtemp(prefix="bpc_bench_")
    path = os.path.join(d, "scene.bscn")
    sc.write(path)
    r = subprocess.run([ORACLE, path, "--steps", str(steps), "--warmup", str(warmup), "--threads", str(threads)],
                       capture_output=True, text=True)
    os.remove(path)
    os.rmdir(d)
    if r.returncode != 0:
        raise SystemExit("oracle_step failed: " + r.stderr[-1000:])
    return json.loads(r.stdout.strip().splitlines()[-1])


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sc, label = build_scene(args.scene, args.scale)
    nd = n_dynamic(sc)
    info = run_reference(args, sc, label, args.steps, args.warmup)
    value = nd * info["steps"] / info["wall_s"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * info["wall_s"] / info["steps"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": label, "bodies": nd, "dt": float(sc.dt), "solver_iterations": sc.iterations},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["threads"], "kind": "reference",
                         "sample": "%d warm-up + %d timed steps of the same scene, WorkerThreadPool with all host threads" % (args.warmup, info["steps"]),
                         "phase_us": info["phase_us"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


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
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc, label = build_scene(args.scene, args.scale, rank, world)
    nd = n_dynamic(sc)
    total_nd = reduce_sum(float(nd))
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 * 1024 * 1024 // 4, dtype=torch.float32, device="cuda")  # 256 MiB > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident arm
    ps = server.PhysicsServerCUDA(local, stream)
    ps.load_scene(sc)
    for _ in range(args.warmup):
        ps.step(sc.dt)
    ps.sync()
    sampler = ClockSampler(local)
    sampler.start()
    ps.set_profiling(True)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    acc = {"contact_pairs": 0, "active_contacts": 0, "collision_pairs": 0, "colors": 0, "launches": 0}
    barrier()
    for i in range(args.steps):
        flush.zero_()
        ev[i][0].record()
        ps.step(sc.dt)
        ev[i][1].record()
        st = ps.stats()  # synchronises; outside the event pair
        for k in ("contact_pairs", "active_contacts", "collision_pairs", "colors"):
            acc[k] += st[k]
        acc["launches"] += st["kernel_launches"]
    barrier()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    process_info = reduce_process_info(st)
    stage_ms, stage_steps = ps.stage_times()
    ps.set_profiling(False)
    sampler.stop_flag = True
    sampler.join()
    ms_max = reduce_max_ms(ms)
    value = total_nd * args.steps / (ms_max * 1e-3)
    ps.finish()

    # ---------------- end-to-end arm: host buffers through the C ABI, fresh context, same schedule
    ps2 = server.PhysicsServerCUDA(local, stream)
    ps2.load_scene(sc)
    nb = ps2.body_count()
    forces_t = torch.zeros((nb, 6), dtype=torch.float32).pin_memory()   # pinned host buffers, as an engine would keep
    xf_t = torch.zeros((nb, 12), dtype=torch.float32).pin_memory()
    forces, xf_host = forces_t.numpy(), xf_t.numpy()
    for _ in range(args.warmup):
        ps2.step(sc.dt)
    ps2.sync()
    barrier()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    checksum = 0.0
    for _ in range(args.steps):
        ps2.bodies_write(server.FIELD_APPLIED_FORCE, forces)     # H2D: per-body force/torque accumulators
        ps2.step(sc.dt)
        xf = ps2.bodies_read(server.FIELD_TRANSFORM, out=xf_host)  # D2H: every body transform (state sync)
        checksum += float(xf[-1, 10])
    e1.record()
    barrier()
    e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3)
    e2e_value = total_nd * args.steps / (reduce_max_ms(e2e_ms) * 1e-3)
    ps2.finish()

    if rank == 0:
        k = max(1, args.steps)
        pc = acc["contact_pairs"] / k
        cbar = acc["active_contacts"] / max(1, acc["contact_pairs"])
        iters = sc.iterations
        # pre-solve 296 + 76 c bytes per pair, each solver iteration 296 + 108 c bytes per pair
        solve_bytes = pc * ((296 + 76 * cbar) + iters * (296 + 108 * cbar))
        if len(sc.joints):  # a joint-iteration: 296 B + 64 B per Jacobian row (pin 3, hinge 6)
            rows = np.where(sc.joints["type"] == 0, 3, 6).astype(np.float64)
            solve_bytes += float(np.sum(iters * (296 + 64 * rows)))
        # the metric's second half: broadphase + narrowphase achieved GB/s (260 B per body, 16 + 664 B per broadphase pair)
        bp_ms = stage_ms["broadphase"] / max(1, stage_steps)
        np_ms = stage_ms["narrowphase"] / max(1, stage_steps)
        bpairs = acc["collision_pairs"] / k
        bp_bytes, np_bytes = 260.0 * nb + 16.0 * bpairs, 664.0 * bpairs
        solve_ms = stage_ms["solve"] / max(1, stage_steps)
        peak, peak_src = peaks()
        traffic, traffic_src = None, None
        try:  # DRAM bytes of one k_solve launch from the committed ncu capture of this scene (profiles/), if there is one
            with open(os.path.join(ROOT, "profiles", "k_solve_traffic.json")) as f:
                t = json.load(f).get(args.scene if args.scale >= 1.0 else "")
            if t:
                traffic, traffic_src = t["traffic_bytes"], "ncu capture at step %d (%d colliding pairs): %s" % (t["step"], t["contact_pairs"], t["capture"])
        except (OSError, ValueError, KeyError):
            pass
        achieved = solve_bytes / (solve_ms * 1e-3) / 1e9 if solve_ms > 0 else 0.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_max / k, "higher_is_better": True, "scaling": "strong" if (args.scene == "c5" and world > 1) else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": label, "bodies": int(total_nd), "dt": float(sc.dt), "solver_iterations": iters,
                       "parallelism": ("spaces sharded over %d ranks, no data-path collective" % world if args.scene == "c5" else
                                       "replicas only (%d independent copies)" % world) if world > 1 else "1 GPU",
                       "l2": "256 MiB buffer written between timed steps (L2 flush)",
                       "avg_broadphase_pairs": acc["collision_pairs"] / k, "avg_contact_pairs": pc, "avg_contacts_per_pair": cbar,
                       "avg_colors": acc["colors"] / k},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(forces.nbytes), "d2h_bytes_per_step": int(nb * 12 * 4),
                    "checksum": checksum},
            "gpu_launches": acc["launches"],
            "process_info": process_info,
            "clocks": sampler.summary(),
            "roofline": {"bound": "hbm", "kernel": "k_solve (pre-solve + %d Gauss-Seidel sweeps, cooperative launch)" % iters,
                         "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "avg_launch_ms": solve_ms, "algorithmic_bytes_per_launch": solve_bytes,
                         "share_of_step": solve_ms / (sum(stage_ms.values()) / max(1, stage_steps)) if stage_steps else None},
            "stage_ms_per_step": {s: v / max(1, stage_steps) for s, v in stage_ms.items()},
            "stage_gbs": {"broadphase": bp_bytes / (bp_ms * 1e-3) / 1e9 if bp_ms > 0 else None,
                          "narrowphase": np_bytes / (np_ms * 1e-3) / 1e9 if np_ms > 0 else None,
                          "broadphase+narrowphase": (bp_bytes + np_bytes) / ((bp_ms + np_ms) * 1e-3) / 1e9 if bp_ms + np_ms > 0 else None,
                          "note": "algorithmic bytes of SURVEY.md 8(d) / stage time (CUDA events), this rank"},
        }
        if args.cpu_baseline:
            kc = min(args.steps, args.cpu_steps)
            info = run_reference(args, sc, label, kc, args.warmup)
            line["cpu_baseline"] = {"value": nd * info["steps"] / info["wall_s"], "unit": UNIT, "cores": info["threads"], "kind": "reference",
                                    "sample": "%d warm-up + %d timed steps of the same scene (oracle/_ref/oracle_step = the reference's own code)"
                                              % (args.warmup, info["steps"]), "phase_us": info["phase_us"]}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=120)
    ap.add_argument("--warmup", type=int, default=60)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--scene", default="c2")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--cpu-steps", type=int, default=20)
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        cuda_arm(args)


if __name__ == "__main__":
    main()
