#!/usr/bin/env python
"""bench.py — the sspp hot path on BASELINE.json's headline workload.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload = "S4a"): synthetic 512^3-voxel map at 0.3 m (seed 42) + 10^6 nodes on a
100^3 lattice at 1.0 m, radius-connect R = 1.5 with the swept-robot collision test, box
(0.5, 0.5, 0.3) -> 27 rays per pair (SURVEY.md §8 d, BASELINE.json configs[3]).

One "step" = one PathPlanner::connectNodes pass over all nodes.
  value  collision-checked edges/s = pairs handed to the collision test / device time, nodes and map
         already resident in HBM (whole pipeline: binning, count, edge test, CSR compaction)
  e2e    the same through the C ABI with HOST buffers (sspp_gpu_connect_nodes_host): pinned node array
         in, CSR out, copies inside the timed region
  roofline  the dominant kernel (k_edges): algorithmic bytes per launch / its CUDA-event duration
  cpu_baseline  the oracle's restated connectNodes (O(N^2), one thread) on a bounded sub-lattice
  visibility    the second half of the metric: occlusion rays/s of the visibility pass
With --impl reference the reference's own CPU connectNodes (oracle/_ref, else the oracle port) is
timed on the host cores on bounded samples of the same scene.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

RES = 0.3
N_VOX = 512
LATTICE = 100
RADIUS = 1.5
BOX = (0.5, 0.5, 0.3)
UNIT = "edges/s"


def metric_name(regions=1):
    """One string for every N (the driver checks that the metric is the same across the scaling runs): weak scaling
    keeps one S4a region — 10^6 nodes in 512^3 voxels — per GPU; config.workload spells out the scene of the run."""
    return "collision-checked edges/s (connectNodes graph build, 1M nodes / 512^3 voxels per GPU)"


# stdout must carry exactly ONE line, the JSON result: file descriptor 1 is pointed at stderr for the
# whole run (NCCL prints its version banner to stdout from C, whatever NCCL_DEBUG_FILE says; so may other
# libraries) and the result line is written to the saved descriptor at the end.
_RESULT_FD = None


def claim_stdout():
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_RESULT_FD if _RESULT_FD is not None else 1, (json.dumps(line) + "\n").encode())


def workload_config(n_gpus, regions=1, assembler=None):
    w = ("S4a: 512^3 voxels @0.3 m (seed 42, 400 boxes + 200 spheres), 10^6 nodes on a 100^3 lattice @1.0 m, "
         "R=1.5, box (0.5,0.5,0.3) -> 27 rays/pair, treat_unknown_as_occupied")
    if regions > 1:
        w = ("S4a x %d (weak scaling): ONE map of 512 x 512 x %d voxels = %d S4a regions stacked along z (seeds 42..%d), ONE search "
             "space of %d x 10^6 nodes (a 100^3 lattice per region), map and node set replicated on every GPU, rank g owns "
             "the source rows of region g; " % (regions, 512 * regions, regions, 41 + regions, regions)) + w
    c = {"workload": w, "nodes": regions * LATTICE ** 3, "voxels": regions * N_VOX ** 3, "radius": RADIUS, "box": list(BOX),
         "parallelism": "rows%d" % n_gpus,
         "l2": "flushed between timed steps (256 MiB write, outside the timed events); per-step scratch (~0.4 GB) also exceeds L2"}
    if assembler:
        c["csr_assembly"] = assembler
    return c


def region_centre_z(g, regions):
    return (g - (regions - 1) / 2.0) * N_VOX * RES


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU legs (oracle = checker; here it is the thing timed as the BASELINE, never the product)
# ------------------------------------------------------------------------------------------------
def sub_scene(occ, known, key_min, lattice_n, centre_idx):
    """A sub-lattice of the S4a scene with a map window just large enough for its reach."""
    from sspp_b200 import scenes
    full_axis = (np.arange(LATTICE) - (LATTICE - 1) / 2.0)
    lo = [int(c - lattice_n // 2) for c in centre_idx]
    ax = [full_axis[lo[a]:lo[a] + lattice_n] for a in range(3)]
    zz, xx, yy = np.meshgrid(ax[2], ax[0], ax[1], indexing="ij")
    nodes = np.ascontiguousarray(np.stack([xx.ravel(), yy.ravel(), zz.ravel()], axis=1)[::-1])
    # voxel window: nodes' bounding box grown by the reach (R + box/2) and two voxels
    half = N_VOX // 2
    w0, w1 = [], []
    for a in range(3):
        w0.append(int(np.floor((ax[a][0] - RADIUS - 0.5) / RES)) - 2 + half)
        w1.append(int(np.floor((ax[a][-1] + RADIUS + 0.5) / RES)) + 3 + half)
    vol_o = scenes.unpack_bits(occ, N_VOX)[w0[2]:w1[2], w0[1]:w1[1], w0[0]:w1[0]]
    vol_k = scenes.unpack_bits(known, N_VOX)[w0[2]:w1[2], w0[1]:w1[1], w0[0]:w1[0]]
    kmin = np.array([key_min[a] + w0[a] for a in range(3)], dtype=np.int32)
    dims = np.array([w1[a] - w0[a] for a in range(3)], dtype=np.int32)
    return nodes, kmin, dims, scenes.pack_bits(vol_o), scenes.pack_bits(vol_k)


def cpu_port_sample(occ, known, key_min, lattice_n=28):
    """cpu_baseline leg: the oracle's restated connectNodes, single thread, on a centred sub-lattice."""
    sys.path.insert(0, os.path.join(REPO, "tests"))
    import oracle_lib
    O = oracle_lib.Oracle()
    nodes, kmin, dims, so, sk = sub_scene(occ, known, key_min, lattice_n, (LATTICE // 2,) * 3)
    w = O.world(RES, True)
    w.build_dense(kmin, dims, so, sk)
    t0 = time.perf_counter()
    rp, col, cnt = O.connect(w, nodes, RADIUS, BOX)
    dt = time.perf_counter() - t0
    w.close()
    return {"value": float(cnt[0]) / dt, "unit": UNIT, "cores": 1, "kind": "port", "tested_pairs": int(cnt[0]),
            "sample": "centred %d^3 sub-lattice of the S4a scene (%d nodes, %d tested pairs, %d voxel lookups), full O(N^2) "
                      "connectNodes incl. the pair scan, %.2f s; at 10^6 nodes the reference's N^2 scan alone is ~1300x larger "
                      "per pair" % (lattice_n, len(nodes), int(cnt[0]), int(cnt[2]), dt),
            "seconds": dt}, (nodes, rp, col)


_REF = {}

VIS = {"nv": 65536, "np": 4_000_000, "hfov": 58.0, "vfov": 45.0, "near": 0.7, "far": 6.0, "standoff": 3.0}
# the far corners of that frustum lie 6 * sqrt(1 + tan^2(29 deg) + tan^2(22.5 deg)) = 7.3 m from the sensor
VIS_POINT_REACH, VIS_MAP_REACH = 7.6, 7.8


def visibility_scene(occ, known, key_min, dims):
    from sspp_b200 import scenes
    pts = scenes.surface_points(occ, known, key_min, dims, RES, VIS["np"], seed=45)
    poses = scenes.viewpoints_near(pts, VIS["nv"], standoff=VIS["standoff"], seed=46)
    return pts, poses


def oracle_visibility_windowed(O, occ, known, key_min, pts, poses):
    """The oracle's visibility (authored spec; frustum test + castRay on a pruned octree) for a few viewpoints,
    each on an octree cut out around it and on the model points within reach of its frustum — a CPU
    implementation with the spatial pruning any sane one has.  -> (view_ptr, point_idx, rays, seconds)"""
    from sspp_b200 import scenes
    vp, idx, rays, dt = [0], [], 0, 0.0
    for v in range(len(poses)):
        c = poses[v, :3]
        near = np.nonzero(np.all(np.abs(pts - c.astype(np.float32)) < VIS_POINT_REACH, axis=1))[0]
        kmin, d, so, sk = scenes.map_window(occ, known, key_min, c - VIS_MAP_REACH, c + VIS_MAP_REACH, RES)
        w = O.world(RES, True)
        w.build_dense(kmin, d, so, sk)
        sub = np.ascontiguousarray(pts[near])
        t0 = time.perf_counter()
        _, idx_o, rays_o, _ = w.visibility(poses[v:v + 1], VIS["hfov"], VIS["vfov"], VIS["near"], VIS["far"], sub)
        dt += time.perf_counter() - t0
        w.close()
        idx.append(near[idx_o])
        vp.append(vp[-1] + len(idx_o))
        rays += int(rays_o)
    return np.array(vp, dtype=np.int64), (np.concatenate(idx) if idx else np.zeros(0, np.int64)), rays, dt


def _vis_worker_init(per_worker, counter):
    sys.path.insert(0, os.path.join(REPO, "tests"))
    import oracle_lib
    from sspp_b200 import scenes
    with counter.get_lock():
        me = counter.value
        counter.value += 1
    O = oracle_lib.Oracle()
    key_min, occ, known = _REF["scene"]
    pts, poses = _REF["vis"]
    sel = poses[me * per_worker:(me + 1) * per_worker]
    jobs = []
    for v in range(len(sel)):
        c = sel[v, :3]
        near = np.nonzero(np.all(np.abs(pts - c.astype(np.float32)) < VIS_POINT_REACH, axis=1))[0]
        kmin, d, so, sk = scenes.map_window(occ, known, key_min, c - VIS_MAP_REACH, c + VIS_MAP_REACH, RES)
        w = O.world(RES, True)
        w.build_dense(kmin, d, so, sk)
        jobs.append((w, sel[v:v + 1], np.ascontiguousarray(pts[near])))
    _REF["vis_jobs"] = jobs


def _vis_worker_step(_):
    t0 = time.perf_counter()
    rays = 0
    for w, pose, sub in _REF["vis_jobs"]:
        _, _, r, _ = w.visibility(pose, VIS["hfov"], VIS["vfov"], VIS["near"], VIS["far"], sub)
        rays += int(r)
    return rays, time.perf_counter() - t0


def reference_visibility_leg(occ, known, key_min, dims, procs, steps, warmup):
    """Visibility half of the metric on the host cores.  The reference snapshot holds no CoverageHeuristic
    source, so what runs is the oracle port of the authored spec, one worker process per core,
    every worker on its own viewpoints of the bench scene with an octree window and the model points in reach."""
    import multiprocessing as mp
    per_worker = 24
    pts, poses = visibility_scene(occ, known, key_min, dims)
    rng = np.random.RandomState(2)
    sel = poses[rng.choice(len(poses), procs * per_worker, replace=False)]
    _REF["vis"] = (pts, sel)
    ctx = mp.get_context("fork")
    counter = ctx.Value("i", 0)
    pool = ctx.Pool(procs, initializer=_vis_worker_init, initargs=(per_worker, counter))
    tot_t, tot_r = 0.0, 0
    for k in range(warmup + steps):
        t0 = time.perf_counter()
        res = pool.map(_vis_worker_step, range(procs), chunksize=1)
        if k >= warmup:
            tot_t += time.perf_counter() - t0
            tot_r += sum(r[0] for r in res)
    pool.close()
    pool.join()
    return {"metric": "visibility rays/s", "value": tot_r / tot_t, "unit": "rays/s", "cores": procs, "kind": "port",
            "sample": "%d worker processes x %d viewpoints of the bench scene's %d (x %d model points) per step; oracle port of the "
                      "authored visibility spec: frustum test over the model points within %.1f m of the sensor + castRay on an "
                      "octree cut out around the viewpoint; %d rays per step" % (procs, per_worker, VIS["nv"], VIS["np"],
                                                                                 VIS_POINT_REACH, tot_r // max(steps, 1))}


def _ref_worker_init(centres, ln, have_ref, counter):
    """One worker process per host thread: its own octree window + the reference's PathPlanner.
    (Processes, not threads: the reference narrates on the process-wide std::cout.)"""
    sys.path.insert(0, os.path.join(REPO, "tests"))
    import oracle_lib
    with counter.get_lock():
        me = counter.value
        counter.value += 1
    O = oracle_lib.Oracle(oracle_lib.LIB_REF if have_ref else oracle_lib.LIB_ORACLE)
    key_min, occ, known = _REF["scene"]
    nodes, kmin, d, so, sk = sub_scene(occ, known, key_min, ln, centres[me % len(centres)])
    w = O.world(RES, True)
    w.build_dense(kmin, d, so, sk)
    p = None
    if have_ref:
        p = O.planner(RADIUS, w)
        poses = np.zeros((len(nodes), 7))
        poses[:, :3] = nodes[::-1]  # insertion order = reverse list order (head insertion)
        poses[:, 6] = 1.0
        p.insert_nodes(poses)
    _REF["job"] = (O, w, p, nodes)


def _ref_worker_step(_):
    O, w, p, nodes = _REF["job"]
    t0 = time.perf_counter()
    if p is not None:
        p.disconnect()
        c0 = p.collision_calls()
        p.connect()  # PathPlanner::connectNodes, src/pathplanner.cpp:527-567, unmodified
        c = p.collision_calls() - c0
    else:
        _, _, cnt = O.connect(w, nodes, RADIUS, BOX)
        c = int(cnt[0])
    return c, time.perf_counter() - t0


def run_reference_arm(args):
    """--impl reference: the reference's own CPU connectNodes on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    sys.path.insert(0, os.path.join(REPO, "tests"))
    import oracle_lib
    from sspp_b200 import scenes
    have_ref = os.path.exists(oracle_lib.LIB_REF)
    # the library is mapped in THIS process before the pool forks, so the workers inherit it and whoever
    # watches this process sees which reference build the arm ran
    _REF["lib"] = oracle_lib.Oracle(oracle_lib.LIB_REF if have_ref else oracle_lib.LIB_ORACLE)
    key_min, dims, occ, known = scenes.synthetic_map(n_vox=N_VOX, res=RES, seed=42)
    _REF["scene"] = (key_min, occ, known)
    cores = os.cpu_count() or 1
    procs = max(1, min(cores, 64))
    ln = 16  # 4096 nodes per worker and step: ~1 s of the reference's single-threaded connectNodes
    rng = np.random.RandomState(1)
    centres = [tuple(int(x) for x in rng.randint(ln, LATTICE - ln, 3)) for _ in range(procs)]
    ctx = mp.get_context("fork")
    counter = ctx.Value("i", 0)
    pool = ctx.Pool(procs, initializer=_ref_worker_init, initargs=(centres, ln, have_ref, counter))

    def one_step():
        t0 = time.perf_counter()
        res = pool.map(_ref_worker_step, range(procs), chunksize=1)
        return time.perf_counter() - t0, sum(r[0] for r in res)

    for _ in range(max(args.warmup, 1)):
        one_step()
    total_t, total_c = 0.0, 0
    for _ in range(args.steps):
        dt, c = one_step()
        total_t += dt
        total_c += c
    pool.close()
    pool.join()
    value = total_c / total_t
    kind = "reference" if have_ref else "port"
    vis = None
    if not args.no_visibility:
        try:
            vis = reference_visibility_leg(occ, known, key_min, dims, procs, max(args.steps, 1), max(args.warmup, 1))
        except Exception as e:  # the headline line must still print
            vis = {"error": repr(e)}
    sample = ("%d worker processes (one per host core) x one %d^3 sub-lattice (%d nodes) of the S4a scene (region 0) per step, each running "
              "%s connectNodes (O(N^2) pair scan + getLineStatusBoundingBox per pair) on its own octree window; %d tested pairs "
              "per step" % (procs, ln, ln ** 3, "the reference's own single-threaded" if have_ref else "the oracle port of",
                            total_c // max(args.steps, 1)))
    line = {"impl": "reference", "metric": metric_name(args.gpus if (args.gpus > 1 and args.scaling == "weak") else 1), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total_t / max(args.steps, 1), "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64+f32 (CPU)", "data": "synthetic",
            "config": workload_config(args.gpus, args.gpus if (args.gpus > 1 and args.scaling == "weak") else 1),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "visibility": vis}
    emit(line)


def profile_traffic(kernel):
    """DRAM / L2 bytes per launch of a kernel from the newest committed `ncu --set full` summary under profiles/."""
    import glob
    import re
    out = {"dram_bytes": None, "lts_bytes": None, "source": None}
    for path in sorted(glob.glob(os.path.join(REPO, "profiles", "r*_%s_ncu_summary.txt" % kernel)), reverse=True):
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        vals = {}
        for ln in open(path):
            m = re.match(r"(dram__bytes_read\.sum|dram__bytes_write\.sum|lts__t_bytes\.sum)\s+([0-9.eE+-]+)\s+(\w+)", ln)
            if m and m.group(3) in unit:
                vals[m.group(1)] = float(m.group(2)) * unit[m.group(3)]
        if "dram__bytes_read.sum" in vals and "dram__bytes_write.sum" in vals:
            out["dram_bytes"] = vals["dram__bytes_read.sum"] + vals["dram__bytes_write.sum"]
            out["lts_bytes"] = vals.get("lts__t_bytes.sum")
            out["source"] = ("%s (ncu --set full of tools/profile_target.py = this scene and these calls, one warm launch per kernel; "
                             "dram__bytes_read.sum + dram__bytes_write.sum, lts__t_bytes.sum)" % os.path.relpath(path, REPO))
            break
    return out


def hbm_peak():
    peaks_path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        return float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def gpu_visibility_leg(eng, torch, ext_stream, occ, known, key_min, dims, K, flush_l2, cpu_leg):
    """Second half of BASELINE's metric, measured like the first: device-timed value (model points resident),
    e2e through the C ABI with host buffers (poses in, index sets out), roofline of the dominant kernel."""
    pts, poses = visibility_scene(occ, known, key_min, dims)
    NV, NP = len(poses), len(pts)
    fr = (VIS["hfov"], VIS["vfov"], VIS["near"], VIS["far"])
    eng.set_model_points(pts)
    st_cnt = eng.visibility(poses, *fr, fetch=False, count_steps=True)  # instrumented: sum of castRay lookups (not timed)
    pin_poses = torch.from_numpy(poses).pin_memory()
    poses_np = pin_poses.numpy()
    for _ in range(3):
        vs = eng.visibility(poses_np, *fr, fetch=False)
    vp_host = torch.empty(NV + 1, dtype=torch.int64).pin_memory()
    idx_host = torch.empty(int(vs.visible_pairs) + 64, dtype=torch.int32).pin_memory()
    launches0 = eng.launch_count()
    tot, phases = 0.0, []
    for _ in range(K):
        flush_l2()
        vs = eng.visibility(poses_np, *fr, fetch=False)
        t = eng.visibility_timings()
        tot += float(t[0])
        phases.append([float(x) for x in t[:5]])
    launches = (eng.launch_count() - launches0) // max(K, 1)
    ph = np.mean(np.array(phases), axis=0)
    ms = tot / K
    e2e_t = 0.0
    for _ in range(K):
        flush_l2()
        t0 = time.perf_counter()
        eng.visibility(poses_np, *fr, vp_out=vp_host.numpy(), idx_out=idx_host.numpy())
        e2e_t += time.perf_counter() - t0
    e2e_ms = e2e_t / K * 1e3
    peak, peak_src = hbm_peak()
    rays, visible, steps = float(vs.rays), float(vs.visible_pairs), float(st_cnt.voxel_steps)
    alg = 12.0 * rays + 4.0 * steps + 4.0 * visible + 28.0 * NV
    k_ms = float(ph[3])
    tr = profile_traffic("k_view_rays")
    try:
        l2_gbps = eng.l2_bandwidth(32 << 20, 30)
    except Exception:
        l2_gbps = None
    achieved = alg / (k_ms * 1e-3) / 1e9
    vis = {"metric": "visibility rays/s", "value": rays / (ms * 1e-3), "unit": "rays/s", "ms_per_pass": ms,
           "timing": "CUDA events on the engine's stream around the whole pass (pose upload, viewpoint set-up, frustum gather, one "
                     "castRay per in-frustum pair, compaction, per-viewpoint sort); model points resident and binned; L2 flushed "
                     "between passes",
           "config": {"workload": "%d viewpoints (3 m from random surface points, seed 46) x %d surface points of the S4a map (seed 45), "
                                  "frustum hFOV 58 vFOV 45 near 0.7 far 6.0" % (NV, NP), "viewpoints": NV, "model_points": NP},
           "rays": int(rays), "visible_pairs": int(visible), "voxel_steps": int(steps), "gpu_launches": int(launches),
           "e2e": {"value": rays / (e2e_ms * 1e-3), "unit": "rays/s", "ms_per_pass": e2e_ms, "h2d_bytes_per_step": NV * 56,
                   "d2h_bytes_per_step": (NV + 1) * 8 + int(visible) * 4,
                   "what": "sspp_gpu_visibility + sspp_gpu_visibility_fetch, pinned host poses in, view_ptr + point_idx out"},
           "roofline": {"bound": "issue (L2-resident map; not a bandwidth-bound kernel)", "kernel": "k_view_rays", "kernel_ms": k_ms,
                        "kernel_share_of_pass": k_ms / ms, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                        "peak_source": peak_src, "algorithmic_bytes": alg,
                        "bytes_model": "12*rays + 4*steps + 4*visible + 28*viewpoints (SURVEY 8 d), rays=%d steps=%d (castRay lookups "
                                       "counted on the device by the instrumented run) visible=%d" % (rays, steps, visible),
                        "l2_read_gbps_measured": l2_gbps, "frac_of_l2": (achieved / l2_gbps) if l2_gbps else None,
                        "traffic": tr["dram_bytes"], "l2_traffic": tr["lts_bytes"], "traffic_source": tr["source"],
                        "measured_dram_gbps": (tr["dram_bytes"] / (k_ms * 1e-3) / 1e9) if tr["dram_bytes"] else None,
                        "measured_l2_gbps": (tr["lts_bytes"] / (k_ms * 1e-3) / 1e9) if tr["lts_bytes"] else None},
           "phase_ms": {"total": float(ph[0]), "binning": float(ph[1]), "gather": float(ph[2]), "rays": float(ph[3]), "compact_sort": float(ph[4])}}
    if cpu_leg:
        sys.path.insert(0, os.path.join(REPO, "tests"))
        import oracle_lib
        O = oracle_lib.Oracle()
        sel = np.sort(np.random.RandomState(3).choice(NV, 256, replace=False))
        vp_o, idx_o, rays_o, dtc = oracle_visibility_windowed(O, occ, known, key_min, pts, poses[sel])
        _, vp_g, idx_g = eng.visibility(poses, *fr)
        same = all(np.array_equal(idx_g[vp_g[v]:vp_g[v + 1]], idx_o[vp_o[k]:vp_o[k + 1]]) for k, v in enumerate(sel))
        vis["cpu_baseline"] = {"value": rays_o / dtc, "unit": "rays/s", "cores": 1, "kind": "port",
                               "sample": "256 viewpoints drawn from all %d; oracle port of the authored spec (frustum test over the "
                                         "model points within %.1f m + castRay on an octree cut out around the viewpoint), %d rays, "
                                         "%.2f s in the oracle" % (NV, VIS_POINT_REACH, rays_o, dtc)}
        vis["parity_on_bench_scene"] = {"viewpoints_checked": int(len(sel)), "rays_checked": int(rays_o),
                                        "visible_checked": int(len(idx_o)), "index_sets_equal": bool(same)}
    return vis


S5 = {"n_vox": 1024, "res": 0.15, "lattice": 216, "spacing": 0.7, "radius": 1.05, "np": 5_000_000, "nv": 262_144}


def s5_leg(torch, dist, local_rank, rank, world, steps=3):
    """BASELINE configs[4], STRONG scaling of one fixed scene over the ranks of this run: 1024^3 voxels
    @0.15 m, 216^3 = 10 077 696 lattice nodes @0.7 m, R 1.05 (100 rays per pair), 5*10^6 model points.
    connect: rows cut by work (sspp_gpu_balanced_cuts), whole CSR assembled on every rank (peer push).
    visibility: viewpoints cut in equal blocks, (view_ptr, point_idx) assembled on every rank the same way.
    The one-GPU time of the same scene is measured in the same run (every rank runs the whole scene alone).
    Times: wall clock between a barrier + device synchronisation on both sides, max over the ranks."""
    from sspp_b200 import scenes, sharded
    from sspp_b200.engine import Engine
    dev = torch.device("cuda", local_rank)
    key_min, dims, occ, known = scenes.synthetic_map(n_vox=S5["n_vox"], res=S5["res"], seed=42)
    nodes = scenes.lattice_nodes(S5["lattice"], S5["spacing"])
    n = len(nodes)
    e = Engine(local_rank)
    e.map_create(S5["res"], key_min, dims, True)
    e.map_upload_dense(occ, known)
    e.set_nodes(nodes)

    def timed(fn, reps):
        out = []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            r = fn()
            torch.cuda.synchronize()
            e.synchronize()
            out.append(time.perf_counter() - t0)
        t = torch.tensor([float(np.mean(out[1:] if len(out) > 1 else out))], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) * 1e3, r

    # ---- connect
    one_ms, st1 = timed(lambda: e.connect(radius=S5["radius"], box=BOX), steps + 1)
    cuts = e.balanced_cuts(world, 8, radius=S5["radius"], box=BOX)
    asm = sharded.PeerCsrAssembler(e, n, dev, cuts=cuts)

    def sharded_connect():
        st = e.connect(radius=S5["radius"], box=BOX, rows=asm.rows)
        asm.assemble(st)
        return st
    sh_ms, stg = timed(sharded_connect, steps + 2)
    dev_ms = torch.tensor([float(e.connect_timings()[0])], dtype=torch.float64, device=dev)
    allms = torch.empty(world, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(allms, dev_ms)
    # equal-count cuts for comparison (what balancing buys)
    asm_eq = sharded.PeerCsrAssembler(e, n, dev)

    def sharded_connect_eq():
        st = e.connect(radius=S5["radius"], box=BOX, rows=asm_eq.rows)
        asm_eq.assemble(st)
        return st
    eq_ms, _ = timed(sharded_connect_eq, steps + 2)
    dev_ms_eq = torch.tensor([float(e.connect_timings()[0])], dtype=torch.float64, device=dev)
    allms_eq = torch.empty(world, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(allms_eq, dev_ms_eq)
    same = True
    if rank == 0:  # the assembled graph against the one-GPU graph of the same scene
        sf = e.connect(radius=S5["radius"], box=BOX)
        rpf, colf = e.connect_fetch(sf)
        rp, col = asm_eq.out_rp.cpu().numpy(), asm_eq.out_col.cpu().numpy()
        same = bool(np.array_equal(rp, rpf) and np.array_equal(col[: sf.edges], colf))
    asm.close()
    asm_eq.close()
    tested = float(st1.tested_pairs)
    out = {"workload": "S5 (BASELINE configs[4]): 1024^3 voxels @0.15 m (seed 42), 216^3 = %d lattice nodes @0.7 m, R 1.05, box "
                       "(0.5,0.5,0.3) -> %d rays/pair; %d model points; strong scaling over %d GPUs, map / nodes / points replicated"
                       % (n, st1.rays_per_pair, S5["np"], world),
           "connect": {"tested_pairs": int(tested), "edges": int(st1.edges), "one_gpu_ms": one_ms, "n_gpu_ms": sh_ms,
                       "speedup": one_ms / sh_ms, "edges_per_s_one_gpu": tested / (one_ms * 1e-3), "edges_per_s_n_gpu": tested / (sh_ms * 1e-3),
                       "row_cuts": "work-balanced (sspp_gpu_balanced_cuts, weight 8 for rows that need the edge test)",
                       "cuts": [int(c) for c in cuts], "connect_device_ms_per_rank": [round(float(x), 3) for x in allms.cpu()],
                       "equal_row_cuts": {"n_gpu_ms": eq_ms, "connect_device_ms_per_rank": [round(float(x), 3) for x in allms_eq.cpu()]},
                       "csr_equals_one_gpu": same,
                       "includes": "binning, count, edge test, CSR, exchange (counts all-gather + peer push + closing all-reduce)"}}
    # ---- visibility
    pts = scenes.surface_points(occ, known, key_min, dims, S5["res"], S5["np"], seed=45)
    poses = scenes.viewpoints_near(pts, S5["nv"], standoff=VIS["standoff"], seed=46)
    fr = (VIS["hfov"], VIS["vfov"], VIS["near"], VIS["far"])
    e.set_model_points(pts)
    v1_ms, vs1 = timed(lambda: e.visibility(poses, *fr, fetch=False), steps + 2)
    vasm = sharded.PeerCsrAssembler(e, len(poses), dev, kind="visibility")
    mine = np.ascontiguousarray(poses[vasm.rows[0]:vasm.rows[1]])

    def sharded_vis():
        vs = e.visibility(mine, *fr, fetch=False)
        vasm.assemble(vs)
        return vs
    vn_ms, vsn = timed(sharded_vis, steps + 2)
    vsame = True
    if rank == 0:
        _, vpf, idxf = e.visibility(poses, *fr)
        vsame = bool(np.array_equal(vasm.out_rp.cpu().numpy(), vpf) and np.array_equal(vasm.out_col.cpu().numpy()[: len(idxf)], idxf))
    vasm.close()
    out["visibility"] = {"viewpoints": int(S5["nv"]), "model_points": int(S5["np"]), "rays": int(vs1.rays), "visible_pairs": int(vs1.visible_pairs),
                         "one_gpu_ms": v1_ms, "n_gpu_ms": vn_ms, "speedup": v1_ms / vn_ms, "rays_per_s_one_gpu": vs1.rays / (v1_ms * 1e-3),
                         "rays_per_s_n_gpu": vs1.rays / (vn_ms * 1e-3), "result_equals_one_gpu": vsame,
                         "note": "%d viewpoints per pass (3 m from random surface points); a pass over all 10^7 way-points of the "
                                 "scene is %d such batches" % (S5["nv"], -(-n // S5["nv"])),
                         "includes": "pose upload, frustum gather, occlusion rays, compaction + sort, exchange of (view_ptr, point_idx)"}
    e.close()
    return out


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="sspp_b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-visibility", action="store_true")
    ap.add_argument("--no-s5", action="store_true", help="N>1: skip the strong-scaling object of BASELINE configs[4] (S5)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N>1: weak = one S4a region per GPU in one map / one search space (default); strong = S4a split over the GPUs")
    ap.add_argument("--assembler", default="peer", choices=["fused", "streamed", "peer", "nccl"],
                    help="N>1: how every rank gets the whole CSR. fused = the edge kernel stores finished rows into all ranks' "
                         "images over NVLink as it goes; streamed = finished row chunks pushed to all ranks while the next chunk is "
                         "computed; peer = one push kernel over NVLink peer memory after the pass (default: the fastest at 2, 4 and "
                         "8 GPUs); nccl = NCCL all-gather + packing")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    from sspp_b200 import scenes
    from sspp_b200.engine import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's version banner / warnings go to stdout by default, which must hold exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n_gpus = world if distributed else 1
    W = max(args.warmup, 3)
    K = args.steps

    # ---- scene (synthetic; the map and the node set are replicated on every rank).  Weak scaling: rank g
    # generates region g (the S4a generator with seed 42+g) and the packed planes are all-gathered, so
    # every rank uploads the same tall map; region g sits at z = region_centre_z(g).
    regions = world if (distributed and args.scaling == "weak") else 1
    key_min, dims, occ, known = scenes.synthetic_map(n_vox=N_VOX, res=RES, seed=42 + (rank if regions > 1 else 0))
    occ0, known0, key_min0, dims0 = occ, known, key_min.copy(), dims.copy()  # this rank's region alone (visibility leg)
    if regions > 1:
        planes = torch.from_numpy(np.stack([occ, known]).view(np.int32)).cuda()
        gathered = torch.empty((regions,) + tuple(planes.shape), dtype=planes.dtype, device="cuda")
        dist.all_gather_into_tensor(gathered, planes)
        allp = gathered.cpu().numpy().view(np.uint32)  # (region, plane, z, y, xw)
        occ = np.ascontiguousarray(np.concatenate([allp[g, 0] for g in range(regions)], axis=0))
        known = np.ascontiguousarray(np.concatenate([allp[g, 1] for g in range(regions)], axis=0))
        del planes, gathered, allp
        key_min = key_min.copy()
        dims = dims.copy()
        key_min[2] = 32768 - (N_VOX // 2) * regions
        dims[2] = N_VOX * regions
        key_min0[2] = key_min[2] + N_VOX * rank
    nodes = np.ascontiguousarray(np.concatenate([scenes.lattice_nodes(LATTICE, 1.0, centre=(0.0, 0.0, region_centre_z(g, regions)))
                                                 for g in range(regions)], axis=0))
    n = len(nodes)
    eng = Engine(local_rank)
    eng.map_create(RES, key_min, dims, True)
    eng.map_upload_dense(occ, known)
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world
    rows = None if not distributed else (lo, hi)
    ext_stream = torch.cuda.ExternalStream(eng.stream(), device=torch.device("cuda", local_rank))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush_l2():
        flush.zero_()
        torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        eng.synchronize()
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- instrumented pass: Σsteps of the byte model (not timed)
    eng.set_nodes(nodes)
    st_cnt = eng.connect(radius=RADIUS, box=BOX, rows=rows, count_steps=True)
    steps_total = st_cnt.voxel_steps

    # ---- device-resident timing: `value`
    assembler, streamed = None, None
    if distributed:
        from sspp_b200 import sharded
        dev = torch.device("cuda", local_rank)
        if args.assembler == "fused":
            # exchange fused into the edge kernel (rows stored into all ranks' slot-format images as they finish)
            streamed = sharded.FusedPeerAssembler(eng, n, dev)
            assembler = sharded.PeerCsrAssembler(eng, n, dev)  # the e2e leg (host buffers in) assembles after its pass
        elif args.assembler == "streamed":
            streamed = sharded.StreamedPeerAssembler(eng, n, dev)
            assembler = sharded.PeerCsrAssembler(eng, n, dev)  # the e2e leg (host buffers in) assembles after its pass
        elif args.assembler == "peer":
            try:
                assembler = sharded.PeerCsrAssembler(eng, n, dev)
            except sharded.PeerMemoryUnavailable as e:  # raised on every rank together
                args.assembler = "nccl (peer memory unavailable: %s)" % e
                assembler = sharded.GpuCsrAssembler(eng, n, dev)
        else:
            assembler = sharded.GpuCsrAssembler(eng, n, dev)

    def assemble_csr(st):
        """N>1: every rank ends the step holding the whole graph's CSR (peer-memory push, or NCCL all-gather + packing)."""
        if assembler is not None:
            return assembler.assemble(st)

    split = []  # per timed step: host seconds in connect / in the CSR assembly (the latter includes waiting for the slowest rank)

    def step_resident():
        t0 = time.perf_counter()
        if streamed is not None:  # connect and exchange in one call
            st, _, _ = streamed.connect(radius=RADIUS, box=BOX)
            split.append((time.perf_counter() - t0, 0.0))
            return st
        st = eng.connect(radius=RADIUS, box=BOX, rows=rows)
        t1 = time.perf_counter()
        assemble_csr(st)
        split.append((t1 - t0, time.perf_counter() - t1))
        return st

    for _ in range(W):
        st = step_resident()
    tested_local = st.tested_pairs
    split.clear()
    flush_l2()
    barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    launches0 = eng.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kern_ms, phase_ms = [], []
    wall = []
    for k in range(K):
        flush_l2()
        if distributed:
            dist.barrier()
        t0 = time.perf_counter()
        ev[k][0].record(ext_stream)
        st = step_resident()
        if distributed:
            torch.cuda.current_stream().synchronize()
        ev[k][1].record(ext_stream)
        eng.synchronize()
        wall.append(time.perf_counter() - t0)
        t = eng.connect_timings()
        kern_ms.append(float(t[5]))
        phase_ms.append([float(x) for x in t])
    barrier()
    launches = eng.launch_count() - launches0
    # device time per step: CUDA events on the engine's stream; the NCCL part (other stream) is
    # covered by the host wall clock of the step, so take the larger of the two when distributed
    step_ms = []
    for k in range(K):
        e = ev[k][0].elapsed_time(ev[k][1])
        step_ms.append(max(e, wall[k] * 1e3) if distributed else e)
    total_ms = float(sum(step_ms))
    tot = torch.tensor([total_ms, float(tested_local), float(st.edges)], dtype=torch.float64, device="cuda")
    if distributed:
        mx = tot.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tot.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_ms = float(mx[0])
        tested_all, edges_all = float(sm[1]), float(sm[2])
    else:
        tested_all, edges_all = float(tested_local), float(st.edges)
    value = tested_all * K / (total_ms * 1e-3)
    multi = None
    if distributed:
        mine = torch.tensor([np.mean([a for a, _ in split]) * 1e3, np.mean([b for _, b in split]) * 1e3,
                             float(np.mean([p[0] for p in phase_ms])), float(tested_local)], dtype=torch.float64, device="cuda")
        allm = torch.empty(world * 4, dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allm, mine)
        allm = allm.cpu().numpy().reshape(world, 4)
        multi = {"connect_host_ms_per_rank": [round(float(x), 4) for x in allm[:, 0]],
                 "assembly_host_ms_per_rank": [round(float(x), 4) for x in allm[:, 1]],
                 "connect_device_ms_per_rank": [round(float(x), 4) for x in allm[:, 2]],
                 "tested_pairs_per_rank": [int(x) for x in allm[:, 3]],
                 "note": "assembly = counts all-gather + push kernel (or NCCL all-gather + packing) + closing barrier, "
                         "host wall clock per rank; it includes the wait for the slowest rank's connect"}

    # ---- e2e: host buffers through the C ABI (pinned), copies inside the timed region
    pin_nodes = torch.from_numpy(nodes).pin_memory()
    rp_host = torch.empty(n + 1, dtype=torch.int64).pin_memory()
    col_host = torch.empty(int(st.in_radius_pairs) + 64, dtype=torch.int32).pin_memory()
    rp_np, col_np, nodes_np = rp_host.numpy(), col_host.numpy(), pin_nodes.numpy()

    if distributed:
        assert n % world == 0
        nodes_dev = torch.empty((n, 3), dtype=torch.float64, device="cuda")

    def step_e2e():
        if not distributed:
            return eng.connect_nodes_host(nodes_np, rp_np, col_np, radius=RADIUS, box=BOX)
        # N > 1: every rank uploads ITS slice of the host node array over its own PCIe link, the slices are
        # all-gathered over NVLink into the replicated node set, the rank connects its rows while its block
        # travels to pinned host memory chunk by chunk, and the whole CSR is assembled on every GPU
        mine = nodes_dev[lo:hi]
        mine.copy_(pin_nodes[lo:hi], non_blocking=True)
        dist.all_gather_into_tensor(nodes_dev, mine)
        ext_stream.wait_event(torch.cuda.current_stream().record_event())
        eng.set_nodes_device(nodes_dev.data_ptr(), n)
        s = eng.connect_to_host(rp_np[: (hi - lo) + 1], col_np, radius=RADIUS, box=BOX, rows=rows)  # block home, chunk by chunk
        assembler.assemble(s)
        return s

    for _ in range(2):
        s2 = step_e2e()
    barrier()
    e2e_t = 0.0
    for k in range(K):
        flush_l2()
        if distributed:
            dist.barrier()
        t0 = time.perf_counter()
        s2 = step_e2e()
        torch.cuda.synchronize()
        e2e_t += time.perf_counter() - t0
    barrier()
    e2e_tot = torch.tensor([e2e_t], dtype=torch.float64, device="cuda")
    if distributed:
        dist.all_reduce(e2e_tot, op=dist.ReduceOp.MAX)
    e2e_value = tested_all * K / float(e2e_tot[0])
    h2d = n * 24  # whole job: every node crosses PCIe once (N > 1: one slice per rank)
    d2h = n * 8 + 8 * n_gpus + int(edges_all) * 4  # whole job: every rank reads its own block
    clk = clocks.stop() if rank == 0 else None

    # ---- correctness guard on the timed configuration: the pyramid skip must not change the CSR
    if rank == 0 and not distributed:
        sa = eng.connect(radius=RADIUS, box=BOX, use_pyramid=False)
        rpa, cola = eng.connect_fetch(sa)
        sb = eng.connect(radius=RADIUS, box=BOX, use_pyramid=True)
        rpb, colb = eng.connect_fetch(sb)
        assert np.array_equal(rpa, rpb) and np.array_equal(cola, colb), "pyramid skip changed the graph"

    # ---- N > 1: strong scaling of the fixed S5 scene (BASELINE configs[4]); every rank takes part, rank 0 reports
    s5, s5_hung = None, False
    if distributed and not args.no_s5:
        # in a thread with a deadline: whatever happens in there (a rank failing inside a collective leaves the others
        # waiting), the headline line above is still printed and every rank still exits
        box = {}

        def run_s5():
            try:
                torch.cuda.set_device(local_rank)
                box["s5"] = s5_leg(torch, dist, local_rank, rank, world)
            except Exception as e:
                box["s5"] = {"error": repr(e)}
        th = threading.Thread(target=run_s5, daemon=True)
        th.start()
        th.join(timeout=float(os.environ.get("SSPP_S5_DEADLINE_S", "420")))
        s5 = box.get("s5", {"error": "did not finish within the deadline"})
        # no collective after this point: a rank that failed or timed out in there must not leave the others waiting
        s5_hung = True

    if rank != 0:
        if distributed and not s5_hung:
            dist.barrier()
            dist.destroy_process_group()
        if s5_hung:
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)
        return

    # ---- roofline of the dominant kernel (k_edges), per launch, this rank's share
    peak, peak_src = hbm_peak()
    C_l, E_l, N_l = float(st.in_radius_pairs), float(st.edges), float(hi - lo)
    alg_bytes = 24.0 * C_l + 4.0 * float(steps_total) + 4.0 * E_l + 32.0 * N_l
    k_ms = float(np.mean(kern_ms))
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    try:
        l2_gbps = eng.l2_bandwidth(32 << 20, 30)
    except Exception:
        l2_gbps = None
    ph = np.mean(np.array(phase_ms), axis=0)
    # `achieved` is the ALGORITHMIC rate (the bytes the reference's work would move, SURVEY 8 d, over the kernel's
    # time): it says how fast the reference's work is disposed of.  What the memory system really carried is
    # measured by ncu per launch of the same command (profiles/): DRAM and L2 bytes, reported beside it.
    tr = profile_traffic("k_edges") if not distributed else {"dram_bytes": None, "lts_bytes": None, "source": None}
    roofline = {"bound": "issue", "bound_note": "the 16 MiB planes are L2-resident and the kernel answers most pairs from key boxes: "
                                                "DRAM and L2 are far from saturated (measured_* below), the limiter ncu names is "
                                                "instruction issue; `achieved`/`frac` follow the contract (algorithmic bytes over the "
                                                "kernel's time against the measured HBM peak), `frac_of_l2` is the north-star's line",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": tr["dram_bytes"], "l2_traffic": tr["lts_bytes"], "traffic_source": tr["source"],
                "measured_dram_gbps": (tr["dram_bytes"] / (k_ms * 1e-3) / 1e9) if tr["dram_bytes"] else None,
                "measured_l2_gbps": (tr["lts_bytes"] / (k_ms * 1e-3) / 1e9) if tr["lts_bytes"] else None,
                "kernel": "k_edges", "kernel_ms": k_ms, "kernel_share_of_step": k_ms / (total_ms / K),
                "algorithmic_bytes": alg_bytes, "peak_source": peak_src,
                "bytes_model": "24*C + 4*steps + 4*E + 32*N with C=%d in-radius pairs, steps=%d voxel lookups (counted on the "
                               "device by the instrumented run, reference semantics), E=%d, N=%d" % (C_l, steps_total, E_l, N_l),
                "l2_read_gbps_measured": l2_gbps, "frac_of_l2": (achieved / l2_gbps) if l2_gbps else None,
                "phase_ms": {"total": float(ph[0]), "binning": float(ph[1]), "count": float(ph[2]), "edge_phase": float(ph[3]),
                             "csr": float(ph[4]), "k_edges": float(ph[5])},
                "pyramid_rows": int(st.pyramid_rows), "box_rows": int(st.tile_rows), "box_pairs": int(st.box_pairs),
                "lattice_pairs": int(st.lattice_pairs)}

    # ---- visibility half of the metric (rank 0, one GPU): occlusion rays/s
    vis = None
    if not args.no_visibility:
        try:
            vis = gpu_visibility_leg(eng, torch, ext_stream, occ0, known0, key_min0, dims0, K, flush_l2,
                                     cpu_leg=(not distributed and not args.no_cpu_baseline))
        except Exception as e:  # the headline line must still print
            vis = {"error": repr(e)}

    # ---- CPU baseline (rank 0, N=1 only), bounded sample
    cpu, parity = None, None
    if not distributed and not args.no_cpu_baseline:
        cpu, (sub_nodes, rp_c, col_c) = cpu_port_sample(occ, known, key_min)
        cpu.pop("seconds", None)
        # parity ON the timed scene: every ordered pair with both ends inside the oracle's sub-lattice must have
        # the same verdict in the GPU's CSR of the WHOLE scene (the oracle saw only the sub-lattice's nodes, the
        # GPU all 10^6: the comparison is on the pairs both know)
        sg = eng.connect(radius=RADIUS, box=BOX)
        rp_g, col_g = eng.connect_fetch(sg)
        ijk = np.rint(sub_nodes + (LATTICE - 1) / 2.0).astype(np.int64)           # lattice coordinates (x, y, z)
        full_idx = n - 1 - ((ijk[:, 2] * LATTICE + ijk[:, 0]) * LATTICE + ijk[:, 1])  # list order = generation order reversed
        assert np.array_equal(nodes[full_idx], sub_nodes)
        inv = np.full(n, -1, dtype=np.int64)
        inv[full_idx] = np.arange(len(sub_nodes))
        rows_g = np.repeat(full_idx, (rp_g[full_idx + 1] - rp_g[full_idx]))
        cols_g = np.concatenate([col_g[rp_g[r]:rp_g[r + 1]] for r in full_idx])
        keep = inv[cols_g] >= 0
        got = np.sort(inv[rows_g[keep]] * len(sub_nodes) + inv[cols_g[keep]])
        rows_c = np.repeat(np.arange(len(sub_nodes)), np.diff(rp_c))
        want = np.sort(rows_c * len(sub_nodes) + col_c)
        bad = np.setxor1d(got, want)
        parity = {"pairs_checked": int(cpu["tested_pairs"]), "oracle_edges": int(len(want)), "gpu_edges_on_those_pairs": int(len(got)),
                  "disagreements": int(len(bad)),
                  "what": "centred 28^3 sub-lattice of the timed scene: oracle connectNodes vs the GPU's whole-scene CSR on every "
                          "ordered pair with both ends in the sub-lattice",
                  "first_disagreeing_pairs": [[int(full_idx[k // len(sub_nodes)]), int(full_idx[k % len(sub_nodes)])] for k in bad[:8]]}
        cpu.pop("tested_pairs", None)

    line = {"metric": metric_name(regions), "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f64 distances + f32/f64 DDA + u32 bit planes", "data": "synthetic",
            "config": workload_config(n_gpus, regions, (args.assembler if distributed else None)),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": float(e2e_tot[0]) * 1e3 / K},
            "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
            "parity_on_bench_scene": parity, "visibility": vis,
            "tested_pairs_per_step": tested_all, "edges_per_step": edges_all, "multi_gpu": multi, "s5": s5}
    emit(line)
    if s5_hung:
        os._exit(0)
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
