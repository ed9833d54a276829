#!/usr/bin/env python
"""bench.py -- the badger_amcl hot path on B200: likelihood-field update + KLD-adaptive systematic resample.

One "step" = PlanarScanner::updateSensor (planar_scanner.cpp:125 -> particle_filter.cpp:223) followed by
ParticleFilter::updateResample (particle_filter.cpp:423) over ONE set of P particles and one B-beam scan on a
W x W occupancy grid (BASELINE.json metric: "particle-beam evals/sec; update+resample ms at 1M particles").
Default workload = north_star's target case: P = 1 000 000 global-uniform particles x 1080 beams, 4000^2 grid,
0.05 m cells, node-default model parameters, min_samples = P/10, max_samples = P.

  value   evals/s with the particle set, map and scan resident in HBM (device-timed, CUDA events on the
          library's stream; the set is restored and L2 flushed between steps, outside the timed region)
  e2e     the same step through the C ABI with HOST buffers: pinned AoS particle upload + scan staging +
          update + resample + AoS download of the resampled set and the result struct, all inside the timer
  roofline   dominant kernel (k_score_lf) alone: algorithmic bytes P*B*4 + P*40 + B*16 / its
             CUDA-event duration vs MEASURED_PEAKS.json hbm_gbs
  cpu_baseline   the reference's own translation units (oracle/_ref, kind "reference") or the C restatement
                 (kind "port") on a bounded particle prefix, one thread (the reference is single-threaded)

`--impl reference` times that CPU path alone. N > 1 (torchrun): every rank runs the step on its own shard of
N*P particles against a replicated map (weak scaling); see DESIGN.md "multi-GPU" for what is exchanged.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RES = 0.05
RANGE_MAX = 30.0
METRIC = "particle-beam evals/sec (likelihood-field update + KLD resample)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--particles", type=int, default=1000000, help="particles per GPU")
    ap.add_argument("--beams", type=int, default=1080)
    ap.add_argument("--map", type=int, default=4000, help="grid is map x map cells")
    ap.add_argument("--cpu-particles", type=int, default=0, help="particle prefix for the CPU baseline (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--legs", default="auto", help="extra BASELINE.json configs measured after the headline "
                    "(comma list of c2,c3,c4,c5,parity; auto = all that apply to --gpus; none)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def workload_name(a):
    return "lf_update+kld_systematic_resample P=%d/gpu B=%d grid=%dx%d res=%.2f" % (a.particles, a.beams, a.map, a.map, RES)


def make_inputs(a, rank=0):
    """Synthetic world of SURVEY 8(d): rooms grid, scan ray-cast from a truth pose, global-uniform particles."""
    from badger_amcl_b200 import synth
    cells = synth.make_grid(a.map, a.map, seed=7)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=a.beams)
    poses = synth.uniform_free_poses(cells, RES, a.particles, seed=1 + rank)
    return cells, truth, ranges, angles, poses


# ------------------------------------------------------------------------------------------------
# clocks sampling (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for nm, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own code (oracle/_ref) or the C restatement (oracle/liboracle.so)
# ------------------------------------------------------------------------------------------------
class CpuArm:
    """updateSensor + updateResample of the reference on the host, one thread (the reference has no threads)."""

    def __init__(self, a, cells, ranges, angles, poses, n_cpu):
        try:
            import resource
            resource.setrlimit(resource.RLIMIT_STACK, (resource.RLIM_INFINITY, resource.RLIM_INFINITY))
        except Exception:
            pass
        from oracle import oracle as O  # bench.py's cpu legs are the one product-side place allowed to run oracle/
        self.O, self.n = O, n_cpu
        self.kind = "reference" if O.have_ref() else "port"
        self.ranges, self.angles = ranges, angles
        self.poses = np.ascontiguousarray(poses[:n_cpu])
        self.model = O.planar_model(model_type=1, max_beams=a.beams, scanner_pose=(0.1, 0.0, 0.0))
        self.valid = int((ranges < RANGE_MAX).sum())
        if self.kind == "reference":
            self.map = O.RefMap2D(cells, RES)
            self.map.update_distances(2.0)  # the reference's own brushfire (setup, untimed)
            self.scanner = O.RefScanner2D(self.map, self.model)
            self.pf = O.RefPF(max(1, n_cpu // 10), n_cpu)
        else:
            self.map = O.PortMap2D(cells, RES)
            self.map.update_distances(2.0)
            self.pf = O.PortPF(max(1, n_cpu // 10), n_cpu)

    def step(self):
        """returns (update_s, resample_s)"""
        self.pf.init_with_poses(self.poses)  # restore the set + its kd-tree (untimed)
        self.pf.srand48(12345)
        t0 = time.perf_counter()
        if self.kind == "reference":
            self.pf.update_sensor(self.scanner, self.ranges, self.angles, RANGE_MAX)
        else:
            self.pf.update_sensor(self.map, self.model, self.ranges, self.angles, RANGE_MAX)
        t1 = time.perf_counter()
        self.pf.update_resample()
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1


def parallel_update_bound(a, arm, cells, threads):
    """SURVEY 8d's "embarrassingly parallel upper bound": the reference has no threads, but its UPDATE is independent per
    particle, so the same object code run by `threads` host threads on disjoint particle shards (each with its own map,
    scanner and filter objects: the classes keep member scratch and are not re-entrant) bounds what a threaded port of
    the update could reach on this host. The resample is one global sequential pass and stays single-threaded.
    Returns the update's wall seconds for arm.n particles, or None."""
    import threading
    O = arm.O
    if arm.kind != "reference" or threads < 2:
        return None
    cache = arm.map.get_distances()
    shards = np.array_split(arm.poses, threads)
    workers = []
    for sh in shards:
        if len(sh) == 0:
            continue
        m = O.RefMap2D(cells, RES)
        m.set_distances(cache, 2.0)
        sc = O.RefScanner2D(m, arm.model)
        pf = O.RefPF(max(1, len(sh) // 10), len(sh))
        pf.init_with_poses(sh)
        workers.append((m, sc, pf))
    gate = threading.Barrier(len(workers) + 1)

    def work(w):
        gate.wait()
        w[2].update_sensor(w[1], arm.ranges, arm.angles, RANGE_MAX)  # ctypes releases the GIL for the call
        gate.wait()

    ts = [threading.Thread(target=work, args=(w,)) for w in workers]
    for t in ts:
        t.start()
    gate.wait()
    t0 = time.perf_counter()
    gate.wait()
    dt = time.perf_counter() - t0
    for t in ts:
        t.join()
    return dt


def run_reference_arm(a):
    rank, _, world = dist_env()
    if rank != 0:
        return 0
    steps, warm = a.steps, a.warmup
    # bounded sample: about 120 s of single-thread work over all steps at ~220 ns per particle-beam
    n_cpu = a.cpu_particles or int(min(50000, max(2000, 120.0 / ((steps + warm) * a.beams * 220e-9))))
    n_cpu = min(n_cpu, a.particles)
    cells, truth, ranges, angles, poses = make_inputs(a)
    arm = CpuArm(a, cells, ranges, angles, poses, n_cpu)
    for _ in range(warm):
        arm.step()
    ts = [arm.step() for _ in range(steps)]
    up, rs = float(np.mean([t[0] for t in ts])), float(np.mean([t[1] for t in ts]))
    value = n_cpu * arm.valid / (up + rs)
    sample = "first %d of %d particles x %d beams (valid %d), update %.1f ms + resample %.1f ms per step" % (
        n_cpu, a.particles, a.beams, arm.valid, up * 1e3, rs * 1e3)
    threads = os.cpu_count() or 1
    bound = None
    try:
        up_par = parallel_update_bound(a, arm, cells, threads)
        if up_par:
            bound = {"value": n_cpu * arm.valid / (up_par + rs), "unit": "evals/s", "threads": threads,
                     "update_ms": up_par * 1e3, "resample_ms": rs * 1e3,
                     "note": "not the reference (it has no threads): its update object code on disjoint particle shards, "
                             "one host thread each; the resample stays sequential"}
    except Exception as e:  # informational leg: never fails the arm
        bound = {"error": str(e)[:200]}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": a.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": (up + rs) * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "timing": "host perf_counter, single thread"},
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": 1, "kind": arm.kind, "sample": sample,
                             "host_cores_available": os.cpu_count(), "parallel_update_upper_bound": bound},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)
    return 0


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_field(kernel, field):
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return float(json.load(f)[kernel][field])
    except Exception:
        return None


def parity_summary():
    """the parity figures scripts/parity_report.py measured on this code (tracked under profiles/), or None"""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_parity.json")) as f:
            return json.load(f)
    except Exception:
        return None


def ncu_traffic(kernel):
    """dram bytes read + written per launch of `kernel`, from the committed ncu --set full capture (profiles/)"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return float(json.load(f)[kernel]["dram_bytes_per_launch"])
    except Exception:
        return None



# ------------------------------------------------------------------------------------------------
# the other BASELINE.json configs: short legs after the headline, reported under config.configs
# ------------------------------------------------------------------------------------------------
def _median_ms(ctx, fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        ctx.flush_l2()
        ctx.sync()
        ctx.timer_start()
        fn()
        ts.append(ctx.timer_stop_ms())
    return float(np.median(ts))


def leg_c2(ctx, cabi, synth, m, cells, ranges, angles, peak):
    """config 2: beam model with ray casting, 100k particles x 1080 beams, 4000^2 grid"""
    P = 100000
    poses = synth.uniform_free_poses(cells, RES, P, seed=1)
    aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
    pf = cabi.ParticleFilter(ctx, P // 10, P)
    model = cabi.planar_model(model_type=cabi.MODEL_BEAM, max_beams=len(ranges), scanner_pose=(0.1, 0.0, 0.0))
    pf.upload_samples(aos)
    pf.stage_scan(m, model, ranges, angles, RANGE_MAX)
    ms = _median_ms(ctx, lambda: pf.update_staged(m, want_total=False))
    k_ms = pf.last_score_ms()
    # Bresenham steps of the same rays (the algorithmic unit of the beam model), from a sample
    sub = poses[:2000]
    steps = 0
    picks = list(range(0, len(ranges), max(1, len(ranges) // 20)))
    for b in picks:
        _, st = m.calc_range(sub[:, 0], sub[:, 1], sub[:, 2] + angles[b], RANGE_MAX, with_steps=True)
        steps += st
    steps_per_ray = steps / (len(sub) * len(picks))
    algo = P * len(ranges) * steps_per_ray * 1.0 + P * 40.0 + len(ranges) * 16.0
    # sampled parity at this shape (untimed): 300 of the particles against the C restatement of calcBeamModel / calcRange
    parity = None
    try:
        from oracle import oracle as O  # checker only
        n = 300
        idx = np.linspace(0, P - 1, n).astype(np.int64)
        sub = np.ascontiguousarray(aos[idx])
        sub[:, 3] = 1.0 / n
        pm = O.PortMap2D(cells, RES)
        pm.set_distances(m.download_distances(), 2.0)
        port = O.PortPF(max(1, n // 10), n)
        port.set_samples(sub)
        port.update_sensor(pm, O.planar_model(model_type=0, max_beams=len(ranges), scanner_pose=(0.1, 0.0, 0.0)), ranges, angles,
                           RANGE_MAX)
        q = cabi.ParticleFilter(ctx, max(1, n // 10), n)
        q.upload_samples(sub)
        q.update_sensor(m, model, ranges, angles, RANGE_MAX)
        got, want = q.weights(), port.samples()[:, 3]
        q.close()
        parity = {"particles": n, "max_rel_weight_diff_vs_port": float(np.abs(got / want - 1.0).max())}
    except Exception as e:  # noqa: BLE001
        parity = {"error": str(e)[:200]}
    pf.close()
    return {"workload": "beam model 100k x %d, 4000^2" % len(ranges), "ms": ms, "kernel": "k_score_beam", "kernel_ms": k_ms,
            "sampled_parity": parity,
            "rays_per_s": P * len(ranges) / (ms * 1e-3), "steps_per_ray": steps_per_ray,
            "cell_steps_per_s": P * len(ranges) * steps_per_ray / (ms * 1e-3), "algorithmic_bytes": algo,
            "achieved_GBps": algo / (k_ms * 1e-3) / 1e9, "frac_of_hbm_peak": algo / (k_ms * 1e-3) / 1e9 / peak}


def leg_c3(ctx, cabi, synth, peak, sizes=(10000, 100000, 1000000, 10000000)):
    """config 3: LF + KLD-adaptive resampling sweep, 2000^2 grid, 1080 beams, min = P/10, max = P, global + tracking"""
    cells = synth.make_grid(2000, 2000, seed=7)
    m = cabi.Map2D(ctx, cells, RES)
    m.build_distances(2.0)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    ranges, angles = synth.make_scan(cells, RES, truth, n_beams=1080)
    valid = int((ranges < RANGE_MAX).sum())
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=1080, scanner_pose=(0.1, 0.0, 0.0))
    rows = []
    for P in sizes:
        for cloud in ("global", "tracking"):
            poses = (synth.uniform_free_poses(cells, RES, P, seed=1) if cloud == "global"
                     else synth.gaussian_poses(truth, P, seed=2))
            aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
            pf = cabi.ParticleFilter(ctx, P // 10, P)
            pf.stage_scan(m, model, ranges, angles, RANGE_MAX)
            up, rs, km = [], [], []
            for rep in range(4):
                pf.upload_samples(aos)
                pf.init_with_current()
                pf.w_slow = pf.w_fast = 0.0
                pf.srand48(1)
                ctx.flush_l2()
                ctx.sync()
                ctx.timer_start()
                pf.update_staged(m, want_total=True)
                up.append(ctx.timer_stop_ms())
                km.append(pf.last_score_ms())
                ctx.timer_start()
                r = pf.update_resample()
                rs.append(ctx.timer_stop_ms())
            u, q, k = float(np.median(up[1:])), float(np.median(rs[1:])), float(np.median(km[1:]))
            algo = float(P) * valid * 4 + P * 40.0 + 1080 * 16.0
            rows.append({"P": P, "cloud": cloud, "update_ms": u, "resample_ms": q, "evals_per_s": P * valid / ((u + q) * 1e-3),
                         "resampled_count": int(r.sample_count), "kld_leaf_count": int(r.leaf_count),
                         "nodes": int(r.node_count), "clusters": int(r.cluster_count),
                         "k_score_lf_frac_of_hbm_peak": algo / (k * 1e-3) / 1e9 / peak,
                         "resample_frac_of_hbm_peak": P * 120.0 / (q * 1e-3) / 1e9 / peak})
            pf.close()
            del aos, poses
    m.close()
    return {"workload": "LF 1080 beams + KLD systematic resample, 2000^2, min=P/10 max=P", "valid_beams": valid, "rows": rows}


def leg_c4(ctx, cabi, synth, peak):
    """config 4: 3-D point-cloud likelihood, 1M particles x 4096 points, 0.05 m voxels, 2000 x 2000 x 60 volume"""
    keys = synth.make_voxel_world(2000, 2000, 60, seed=3)
    m3 = cabi.Map3D(ctx, RES, keys, 0.36)
    ctx.timer_start()
    m3.build_distances()
    edt_ms = ctx.timer_stop_ms()
    truth3 = (1.3, -0.7, 0.4)
    pts = synth.make_cloud(keys, RES, truth3, n_points=4096)
    P = 1000000
    cells2 = synth.make_grid(2000, 2000, seed=7)
    out = {"workload": "3-D cloud likelihood 1M x %d points, %s voxels" % (len(pts), "x".join(str(v) for v in m3.shape())),
           "edt_build_ms": edt_ms}
    cm = cabi.cloud_model(max_beams=4096, mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0, 0.1, 0))
    for cloud in ("global", "tracking"):
        poses = (synth.uniform_free_poses(cells2, RES, P, seed=1) if cloud == "global"
                 else synth.gaussian_poses(truth3, P, seed=2))
        aos = np.concatenate([poses, np.full((P, 1), 1.0 / P)], axis=1)
        pf = cabi.ParticleFilter(ctx, P // 10, P)
        pf.upload_samples(aos)
        pf.stage_cloud(m3, cm, pts)
        ms = _median_ms(ctx, lambda: pf.update_cloud_staged(m3, want_total=False), reps=4, warm=1)
        k_ms = pf.last_score_ms()
        algo = float(P) * len(pts) + P * 40.0 + len(pts) * 12.0
        out[cloud] = {"ms": ms, "kernel": "k_score_cloud", "kernel_ms": k_ms, "evals_per_s": P * len(pts) / (ms * 1e-3),
                      "algorithmic_bytes": algo, "achieved_GBps": algo / (k_ms * 1e-3) / 1e9,
                      "frac_of_hbm_peak": algo / (k_ms * 1e-3) / 1e9 / peak}
        if cloud == "global":
            # sampled parity at this shape (untimed): 64 of the particles against the C restatement over the same field
            try:
                from oracle import oracle as O  # checker only
                n = 64
                idx = np.linspace(0, P - 1, n).astype(np.int64)
                sub = np.ascontiguousarray(aos[idx])
                pm3 = O.PortMap3D(RES, keys, 0.36)
                pm3.set_dense(m3.download_distances())
                ocm = O.cloud_model(max_beams=4096, mount_translation=(0.2, 0.0, 1.5), mount_rotation=synth.rpy_quaternion(0, 0.1, 0))
                want_total, want = pm3.cloud_apply(ocm, pts, sub)
                q = cabi.ParticleFilter(ctx, 8, n)
                q.upload_samples(sub)
                q.score_cloud(m3, cm, pts)
                got = q.weights()
                q.close()
                out["sampled_parity"] = {"particles": n, "max_rel_weight_diff_vs_port": float(np.abs(got / want - 1.0).max())}
            except Exception as e:  # noqa: BLE001
                out["sampled_parity"] = {"error": str(e)[:200]}
        pf.close()
    m3.close()
    return out


def leg_c5(ctx, cabi, sharded, comm, dist, m, model, cells, ranges, angles, rank, world, per_gpu=8000000, steps=3, warmup=1):
    """config 5: global localisation, 64M particles x 1080 beams over 8 GPUs (8M per GPU): one global set, every rank
    generates and owns its window (global-uniform free-space poses); the same step as the headline"""
    import torch
    from badger_amcl_b200 import synth
    PT = per_gpu * world
    off, cnt = (PT * rank) // world, (PT * (rank + 1)) // world - (PT * rank) // world
    poses = synth.uniform_free_poses(cells, RES, cnt, seed=1000 + rank)
    aos = cabi.pinned_array((cnt, 4))
    aos[:, :3] = poses
    aos[:, 3] = 1.0 / PT
    del poses
    pf = cabi.ParticleFilter(ctx, PT // 10, PT)
    sf = sharded.ShardedFilter(pf, comm, m)
    pf.stage_scan(m, model, ranges, angles, RANGE_MAX)
    valid = int((ranges < RANGE_MAX).sum())
    ups, rss = [], []
    r = None
    for it in range(warmup + steps):
        cabi._check(pf.L.bamcl_pf_shard_set_count(pf.h, PT, None, None))
        sf.upload_window(aos)
        sf.init_stats()
        sf.srand48(12345)
        ctx.flush_l2()
        ctx.sync()
        dist.barrier()
        ctx.timer_start()
        sf.update_sensor()
        u = ctx.timer_stop_ms()
        ctx.timer_start()
        r = sf.update_resample()
        q = ctx.timer_stop_ms()
        if it >= warmup:
            ups.append(u)
            rss.append(q)
    t = torch.tensor([float(np.mean(ups)), float(np.mean(rss))], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    up, rs = [float(v) for v in t.tolist()]
    pf.close()
    return {"workload": "LF update + KLD systematic resample, ONE global set of %d particles x %d beams over %d GPUs" % (PT, len(ranges), world),
            "global_particles": PT, "ms_per_step": up + rs, "update_ms": up, "resample_ms": rs,
            "evals_per_s": float(PT) * valid / ((up + rs) * 1e-3), "resampled_count": int(r.sample_count),
            "kld_leaf_count": int(r.leaf_count), "clusters": int(r.cluster_count), "steps": steps}


def leg_host_api(a, cells, ranges, angles, poses, steps=5, warmup=2):
    """the same step through the C++ class API (libbadger_amcl.so: OccupancyMap, PlanarScanner::updateSensor,
    ParticleFilter::updateResample, node_2d.cpp:367-380) with a host-side write of the samples between steps; the
    standalone C++ driver badger_amcl_b200/host/test/test_host_api (mode `bench`) does the timing"""
    import struct
    import tempfile
    exe = os.path.join(ROOT, "badger_amcl_b200", "host", "test", "test_host_api")
    if not os.path.exists(exe):
        return {"error": "host driver not built (python -c 'import __graft_entry__ as g; g.build()')"}
    n = len(poses)
    with tempfile.NamedTemporaryFile(suffix=".bin", delete=False) as f:
        f.write(struct.pack("<9i", 0x414D434C, cells.shape[1], cells.shape[0], n, len(ranges), n // 10, 1, 1, 12345))
        f.write(struct.pack("<3d", RES, 2.0, RANGE_MAX))
        f.write(np.ascontiguousarray(cells, np.int8).tobytes())
        f.write(np.ascontiguousarray(ranges, np.float64).tobytes())
        f.write(np.ascontiguousarray(angles, np.float64).tobytes())
        f.write(np.ascontiguousarray(poses[:, :3], np.float64).tobytes())
        path = f.name
    try:
        r = subprocess.run([exe, "bench", path, str(steps), str(warmup)], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            return {"error": "rc %d: %s" % (r.returncode, (r.stdout + r.stderr)[-400:])}
        return json.loads(r.stdout.strip().splitlines()[-1])
    finally:
        os.unlink(path)


def leg_parity(cabi, pf, m, model, cells, ranges, angles, aos, P):
    """20k particles of the bench set scored on the device against the C restatement of the reference (the checker;
    not timed, after every measurement): max relative difference of the log weight factors, default and exact mode"""
    from oracle import oracle as O  # checker only
    n = min(20000, P)
    idx = np.linspace(0, P - 1, n).astype(np.int64)
    sub = np.ascontiguousarray(aos[idx])
    pm = O.PortMap2D(cells, RES)
    pm.set_distances(m.download_distances(), 2.0)  # the same cache on both sides: this isolates the scorer
    port = O.PortPF(max(1, n // 10), n)
    port.set_samples(sub)
    port.update_sensor(pm, O.planar_model(model_type=1, max_beams=len(ranges), scanner_pose=(0.1, 0.0, 0.0)), ranges, angles,
                       RANGE_MAX)
    want = port.samples()[:, 3]
    out = {"particles": n, "beams": int(len(ranges)), "against": "oracle port (C restatement), same distance cache"}
    q = cabi.ParticleFilter(pf.ctx, max(1, n // 10), n)
    for name, exact in (("default_fixed_point", False), ("exact_terms", True)):
        q.upload_samples(sub)
        q.set_exact_terms(exact)
        q.update_sensor(m, model, ranges, angles, RANGE_MAX)
        got = q.weights()
        # normalised weight = factor / total: compare the logs of the weights (the totals cancel to 1e-16)
        rel = np.abs(np.log(got) - np.log(want)) / np.maximum(np.abs(np.log(want)), 1e-300)
        out[name] = {"max_rel_log_weight_diff": float(rel.max()), "max_rel_weight_diff": float(np.abs(got / want - 1.0).max())}
    q.close()
    return out


def run_b200_arm(a):
    rank, local_rank, world = dist_env()
    from badger_amcl_b200 import cabi
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = cabi.Context(local_rank)
    # ONE global set of world * P particles (weak scaling: P per GPU); map and scan replicated on every rank
    P, B = a.particles, a.beams
    PT = P * world
    ga = argparse.Namespace(**vars(a))
    ga.particles = PT
    cells, truth, ranges, angles, poses = make_inputs(ga)
    valid = int((ranges < RANGE_MAX).sum())
    m = cabi.Map2D(ctx, cells, RES)
    ctx.timer_start()
    m.build_distances(2.0)
    edt_ms = ctx.timer_stop_ms()
    model = cabi.planar_model(model_type=cabi.MODEL_LF, max_beams=B, scanner_pose=(0.1, 0.0, 0.0))
    pf = cabi.ParticleFilter(ctx, PT // 10, PT)
    aos = cabi.pinned_array((PT, 4))
    aos[:, :3] = poses
    aos[:, 3] = 1.0 / PT
    out_aos = cabi.pinned_array((PT, 4))
    pf.srand48(12345)
    sf = comm = None
    if world > 1:
        from badger_amcl_b200 import sharded
        comm = sharded.Comm.nccl(ctx, rank, world, dist)  # the library's own NCCL communicator (data path)
        sf = sharded.ShardedFilter(pf, comm, m)
    off0, cnt0 = (PT * rank) // world, (PT * (rank + 1)) // world - (PT * rank) // world

    def restore():
        # every step resamples the SAME set with the SAME draw, so runs with different --steps / --gpus are comparable
        if sf is None:
            pf.upload_samples(aos)
            pf.init_with_current()  # kd-tree of the set being resampled (particle_filter.cpp:292 reads its leaf count)
            pf.w_slow = pf.w_fast = 0.0
            pf.srand48(12345)
        else:
            cabi._check(pf.L.bamcl_pf_shard_set_count(pf.h, PT, None, None))
            sf.upload_window(aos[off0:off0 + cnt0])  # each rank holds only the window it owns
            sf.init_stats()
            sf.srand48(12345)

    def barrier():
        ctx.sync()
        if dist is not None:
            dist.barrier()

    # ---- resident: scan staged once, set restored outside the timed region
    pf.stage_scan(m, model, ranges, angles, RANGE_MAX)

    def step_resident():
        restore()
        ctx.flush_l2()
        barrier()
        ctx.timer_start()
        if sf is None:
            pf.update_staged(m, want_total=True)
        else:
            sf.update_sensor()       # score this rank's window, all-gather the totals, normalise
        t_up = ctx.timer_stop_ms()
        k_ms = pf.last_score_ms()
        ctx.timer_start()
        r = pf.update_resample() if sf is None else sf.update_resample()
        t_rs = ctx.timer_stop_ms()
        return t_up, t_rs, k_ms, r

    e2e_split = []  # wall ms of the update half of every single-GPU e2e step

    def step_e2e():
        restore()
        ctx.flush_l2()
        barrier()
        t0 = time.perf_counter()
        ctx.timer_start()
        if sf is None:
            # H2D: P x 32 B AoS from pinned host memory (chunks, scored while the next chunk is on the bus) + the scan;
            # D2H: total
            pf.update_sensor_host(m, model, ranges, angles, RANGE_MAX, aos)
            e2e_split.append((time.perf_counter() - t0) * 1e3)  # the update call returns with the total on the host
            # D2H: result struct + n x 32 B AoS of the new set (copied while the pose-bin tree is built)
            r = pf.update_resample(out=out_aos)
            n = r.sample_count
            moved = n
        else:
            # every rank moves ITS window: upload P x 32 B, stage the scan, sharded update + resample, download its
            # share of the new set
            sf.upload_window(aos[off0:off0 + cnt0])
            pf.stage_scan(m, model, ranges, angles, RANGE_MAX)
            sf.update_sensor()
            r = sf.update_resample()
            n = r.sample_count
            o, c = sf.window()
            sf.download_window(out_aos[o:o + c])
            moved = c
        ms = ctx.timer_stop_ms()
        wall = (time.perf_counter() - t0) * 1e3
        return max(ms, wall), moved

    def result_hash(r):
        """64-bit digest of (resample indices, KLD leaf count, sample count) of the last resample, independent of how
        the set is split over ranks: sum over output slots of mix(slot, source) mod 2^64 (+ the two counts)"""
        if sf is None:
            o, src = 0, pf.resample_sources().astype(np.uint64)
        else:
            o, c = sf.window()
            src = sf.window_sources().astype(np.uint64)
        with np.errstate(over="ignore"):
            z = (np.arange(o, o + len(src), dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15)) ^ (src + np.uint64(1))
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            h = int((z ^ (z >> np.uint64(31))).sum(dtype=np.uint64))
        if dist is not None:
            import torch
            t = torch.tensor([h - (1 << 64) if h >= (1 << 63) else h], dtype=torch.int64, device="cuda")
            dist.all_reduce(t)
            h = int(t.item()) & ((1 << 64) - 1)
        h = (h + 0x100000001B3 * int(r.leaf_count) + 0x1000193 * int(r.sample_count)) & ((1 << 64) - 1)
        return "%016x" % h

    for _ in range(a.warmup):
        step_resident()
    launches0 = ctx.launch_count()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    rows = [step_resident() for _ in range(a.steps)]
    barrier()
    launches_timed = ctx.launch_count() - launches0
    up = float(np.sum([r[0] for r in rows]))
    rs = float(np.sum([r[1] for r in rows]))
    k_ms = float(np.mean([r[2] for r in rows]))
    last = rows[-1][3]
    digest = result_hash(last)
    for _ in range(max(1, a.warmup // 2)):
        step_e2e()
    barrier()
    e_rows = [step_e2e() for _ in range(a.steps)]
    barrier()
    clocks = sampler.stop()
    # the step before the path, reported beside the metric, not inside it
    motion_ms = None
    if sf is None:
        pf.srand48(1)
        ts = []
        for _ in range(4):
            ctx.sync()
            ctx.timer_start()
            pf.update_action(cabi.ODOM_DIFF_CORRECTED, [0.2, 0.2, 0.2, 0.2, 0.2], [1.0, 2.0, 0.7], [0.3, 0.1, 0.2])
            ts.append(ctx.timer_stop_ms())
        motion_ms = float(np.median(ts[1:]))
    e_ms = float(np.sum([r[0] for r in e_rows]))
    n_out = int(e_rows[-1][1])

    c5 = None
    want_c5 = (a.legs == "auto" and world == 8 and a.particles == 1000000) or (a.legs != "auto" and "c5" in a.legs.split(","))
    if want_c5 and sf is not None:
        per = 8000000 if world == 8 else max(1000000, 8000000 // 4)
        try:
            c5 = leg_c5(ctx, cabi, sharded, comm, dist, m, model, cells, ranges, angles, rank, world, per_gpu=per)
        except Exception as e:  # noqa: BLE001  (a leg must not take the headline line down)
            c5 = {"error": str(e)[:300]}
    total_ms, e_total_ms = up + rs, e_ms
    if dist is not None:
        import torch
        t = torch.tensor([total_ms, e_total_ms, up, rs, k_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e_total_ms, up, rs, k_ms = [float(v) for v in t.tolist()]
    if rank != 0:
        if comm is not None:
            comm.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    evals = float(PT) * valid
    value = evals * a.steps / (total_ms * 1e-3)
    e_value = evals * a.steps / (e_total_ms * 1e-3)
    peak, peak_src = measured_peaks()
    algo_bytes = float(cnt0) * valid * 4 + float(cnt0) * 40 + B * 16   # one launch scores this rank's window
    achieved = algo_bytes / (k_ms * 1e-3) / 1e9
    par = ("single GPU" if world == 1 else
           "ONE global set of %d particles owned rank-major by %d ranks; map+scan replicated; NCCL inside the C library: "
           "weight totals all-gathered, CDF tile summaries all-gathered (exact rank offsets, no chain), selected poses to "
           "slot owners by grouped send/recv (all-to-all-v), pose-bin first occurrences all-reduced (min) on a dense grid, "
           "bucket subtrees of the KLD tree split over ranks, cluster sums + counts all-reduced" % (PT, world))
    line = {"metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "valid_beams": valid, "model": "likelihood_field",
                       "global_particles": PT,
                       "resample": "kld systematic, min=P/10 max=P, kld_err 0.01 kld_z 0.99",
                       "parallelism": par,
                       "l2": "flushed (256 MiB write) before every timed step",
                       "update_ms": up / a.steps, "resample_ms": rs / a.steps, "edt_build_ms": edt_ms,
                       "resampled_count": int(last.sample_count), "kld_leaf_count": int(last.leaf_count),
                       "clusters": int(last.cluster_count), "result_hash": digest,
                       "motion_update_ms_not_in_metric": motion_ms},
            "e2e": {"value": e_value, "unit": "evals/s", "ms_per_step": e_total_ms / a.steps,
                    "h2d_bytes_per_step": int(cnt0 * 32 + 2 * B * 8) * world,
                    "d2h_bytes_per_step": int(n_out * 32 + 8 + 96) * world,
                    "update_ms": (float(np.mean(e2e_split[-a.steps:])) if e2e_split else None)},
            "gpu_launches": int(launches_timed),
            "roofline": {"kernel": "k_score_lf", "bound": "hbm", "bound_note": "HBM fraction as the contract asks; the kernel's "
                         "field is L2-resident and its real roof is the L2 gather rate, see l2_gather", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": ncu_traffic("k_score_lf") if world == 1 else None,
                         "kernel_ms": k_ms, "peak_source": peak_src, "algorithmic_bytes": algo_bytes},
            "clocks": clocks}
    # ---- what bounds the dominant kernel besides HBM: the rank field is L2-resident, so the roof that matters is the
    # L2 random-gather rate (measured here with the scorer's access pattern stripped of its arithmetic)
    try:
        g_l2 = ctx.gather_rate(32 << 20, 1 << 30)
        line["roofline"]["l2_gather"] = {
            "what": "random 2-byte gathers over a 32 MB (L2-resident) table, 16 in flight per thread; one 32-B sector each",
            "peak_gathers_per_s": g_l2, "peak_sector_GBps": g_l2 * 32 / 1e9,
            "k_score_lf_gathers_per_s": float(cnt0) * valid / (k_ms * 1e-3),
            "k_score_lf_sectors_per_32_lane_request": ncu_field("k_score_lf", "sectors_per_request"),
            "note": "the kernel issues one gather per eval; lanes of a request share sectors (see sectors_per_request), so "
                    "its sector rate = gathers/s x sectors_per_request / 32"}
        spr = ncu_field("k_score_lf", "sectors_per_request")
        if spr:
            line["roofline"]["l2_gather"]["k_score_lf_sector_rate_frac_of_peak"] = (
                float(cnt0) * valid / (k_ms * 1e-3) * spr / 32.0 / g_l2)
    except Exception as e:  # noqa: BLE001
        line["roofline"]["l2_gather"] = {"error": str(e)}
    if c5 is not None:
        line["config"].setdefault("configs", {})["c5_global_64M"] = c5
    line["config"]["distance_cache"] = ("device-built exact EDT (OccupancyMap::updateDistancesLUT on the device); the reference's "
                                        "heap-order brushfire equals it in 99.1 % of this map's cells and is never smaller -- "
                                        "figures and the effect on the weights: config.parity_report / profiles/r02_parity.json")
    legs = [] if a.legs == "none" else (["c2", "c3", "c4", "parity", "host"] if a.legs == "auto" else a.legs.split(","))
    if world == 1 and legs:
        from badger_amcl_b200 import synth
        cfgs = {}
        if "parity" in legs:
            cfgs["sampled_parity_at_bench_shape"] = leg_parity(cabi, pf, m, model, cells, ranges, angles, aos, P)
            line["config"]["parity_report"] = parity_summary()
        pf.close()
        if "c2" in legs:
            cfgs["c2_beam_model"] = leg_c2(ctx, cabi, synth, m, cells, ranges, angles, peak)
        m.close()
        if "host" in legs:
            cfgs["host_cpp_class_api"] = leg_host_api(a, cells, ranges, angles, poses)
        if "c3" in legs:
            cfgs["c3_lf_kld_sweep"] = leg_c3(ctx, cabi, synth, peak)
        if "c4" in legs:
            cfgs["c4_cloud_3d"] = leg_c4(ctx, cabi, synth, peak)
        line["config"].setdefault("configs", {}).update(cfgs)
        line["config"]["configs_note"] = ("config 1 (5k x 60, CPU-runnable) is the parity-test case of tests/test_golden.py; config "
                                          "5 (64M over 8 GPUs) is measured by `--gpus 8 --particles 8000000`")
    if not a.no_cpu_baseline and world == 1:
        n_cpu = a.cpu_particles or min(P, 20000)
        arm = CpuArm(a, cells, ranges, angles, poses, n_cpu)
        arm.step()
        cu, cr = arm.step()
        line["cpu_baseline"] = {"value": n_cpu * arm.valid / (cu + cr), "unit": "evals/s", "cores": 1, "kind": arm.kind,
                                "sample": "first %d of %d particles x %d beams: update %.1f ms + resample %.1f ms, 1 thread"
                                          % (n_cpu, P, B, cu * 1e3, cr * 1e3),
                                "host_cores_available": os.cpu_count()}
    emit(line)
    if comm is not None:
        comm.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def emit(line):
    """the ONE JSON line goes to the real stdout; everything else this process (or NCCL) prints went to stderr"""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    a = parse_args()
    # NCCL and friends write banners to fd 1 ("NCCL version ..."): keep stdout for the JSON line only
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        return run_reference_arm(a)
    return run_b200_arm(a)


if __name__ == "__main__":
    sys.exit(main())
