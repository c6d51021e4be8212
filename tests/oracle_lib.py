"""ctypes binding of the ORACLE (oracle/liboracle.so, oracle/_ref/libsspp_ref.so).  Test infrastructure:
only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this."""
import ctypes
import os

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_ORACLE = os.path.join(REPO, "oracle", "liboracle.so")
LIB_REF = os.path.join(REPO, "oracle", "_ref", "libsspp_ref.so")
LIB_PROBE = os.path.join(REPO, "tests", "cpu_probe", "libdda_probe.so")

vp = ctypes.c_void_p
lng = ctypes.c_long
dbl = ctypes.c_double


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def _f64(a, cols=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a if cols is None else a.reshape(-1, cols)


class World:
    def __init__(self, orc, resolution, treat_unknown):
        self.o = orc
        self.L = orc.L
        self.res = resolution
        self.h = self.L.orc_world_create(resolution, int(bool(treat_unknown)))

    def close(self):
        if self.h:
            self.L.orc_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_box(self, pos, box, occupied):
        self.L.orc_world_set_box(self.h, _p(_f64(pos)), _p(_f64(box)), int(bool(occupied)))

    def set_keys(self, keys, occupied):
        k = np.ascontiguousarray(keys, dtype=np.int32).reshape(-1, 3)
        self.L.orc_world_set_keys(self.h, _p(k), len(k), int(bool(occupied)))

    def insert_pointcloud(self, pts, origin, max_range):
        a = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        self.L.orc_world_insert_pointcloud(self.h, _p(a), len(a), _p(_f64(origin)), float(max_range))

    def export_logodds(self, key_min, dims):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        out = np.zeros((int(dm[2]), int(dm[1]), int(dm[0])), dtype=np.float32)
        self.L.orc_world_export_logodds(self.h, _p(km), _p(dm), _p(out))
        return out

    def prune(self):
        self.L.orc_world_prune(self.h)

    def num_tree_nodes(self):
        return self.L.orc_world_num_tree_nodes(self.h)

    def build_dense(self, key_min, dims, occ_bits, known_bits, prune=True):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        o = np.ascontiguousarray(occ_bits, dtype=np.uint32)
        k = np.ascontiguousarray(known_bits, dtype=np.uint32)
        self.L.orc_world_build_dense(self.h, _p(km), _p(dm), _p(o), _p(k), int(bool(prune)))

    def export_boxes(self, occupied):
        n = self.L.orc_world_export_boxes(self.h, int(bool(occupied)), None, None, 0)
        c = np.zeros((max(n, 1), 3))
        s = np.zeros(max(n, 1))
        self.L.orc_world_export_boxes(self.h, int(bool(occupied)), _p(c), _p(s), n)
        return c[:n], s[:n]

    def export_dense(self, key_min, dims):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        wpr = (int(dm[0]) + 31) // 32
        o = np.zeros((int(dm[2]), int(dm[1]), wpr), dtype=np.uint32)
        k = np.zeros_like(o)
        self.L.orc_world_export_dense(self.h, _p(km), _p(dm), _p(o), _p(k))
        return o, k

    def ray_keys(self, origin, end, cap=4096):
        o = np.asarray(origin, dtype=np.float32)
        e = np.asarray(end, dtype=np.float32)
        keys = np.zeros((cap, 3), dtype=np.int32)
        n = self.L.orc_ray_keys(self.h, _p(o), _p(e), _p(keys), cap)
        return n, keys[:max(min(n, cap), 0)]

    def line_status(self, starts, ends):
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        out = np.zeros(len(s), dtype=np.uint8)
        self.L.orc_line_status(self.h, _p(s), _p(e), len(s), _p(out))
        return out

    def line_status_bbox(self, starts, ends, box, with_counts=False):
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        out = np.zeros(len(s), dtype=np.uint8)
        steps, rays = lng(), lng()
        self.L.orc_line_status_bbox(self.h, _p(s), _p(e), len(s), _p(_f64(box)), _p(out), ctypes.byref(steps), ctypes.byref(rays))
        return (out, steps.value, rays.value) if with_counts else out

    def line_status_bbox_mt(self, starts, ends, box, threads):
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        out = np.zeros(len(s), dtype=np.uint8)
        self.L.orc_line_status_bbox_mt(self.h, _p(s), _p(e), len(s), _p(_f64(box)), _p(out), threads)
        return out

    def cell_status_point(self, pts):
        a = _f64(pts, 3)
        out = np.zeros(len(a), dtype=np.uint8)
        self.L.orc_cell_status_point(self.h, _p(a), len(a), _p(out))
        return out

    def cell_status_bbox(self, pts, box):
        a = _f64(pts, 3)
        out = np.zeros(len(a), dtype=np.uint8)
        self.L.orc_cell_status_bbox(self.h, _p(a), len(a), _p(_f64(box)), _p(out))
        return out

    def cast_rays(self, origins, dirs, ignore_unknown=True, max_range=None):
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        mr = None if max_range is None else _f64(max_range)
        hit = np.zeros(len(o), dtype=np.uint8)
        keys = np.zeros((len(o), 3), dtype=np.int32)
        ends = np.zeros((len(o), 3), dtype=np.float32)
        self.L.orc_cast_rays(self.h, _p(o), _p(d), len(o), int(bool(ignore_unknown)), _p(mr), _p(hit), _p(keys), _p(ends))
        return hit, keys

    def visibility(self, poses, hfov, vfov, near, far, pts):
        P = _f64(poses, 7)
        pt = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        fr = self.o.frustum_consts(hfov, vfov, near, far)
        vptr = np.zeros(len(P) + 1, dtype=np.int64)
        cnt = np.zeros(2, dtype=np.int64)
        n = self.L.orc_visibility(self.h, _p(P), len(P), _p(fr), _p(pt), len(pt), _p(vptr), None, 0, _p(cnt))
        idx = np.zeros(max(n, 1), dtype=np.int32)
        self.L.orc_visibility(self.h, _p(P), len(P), _p(fr), _p(pt), len(pt), _p(vptr), _p(idx), n, _p(cnt))
        return vptr, idx[:n], int(cnt[0]), int(cnt[1])


class RefPlanner:
    """The reference's own PathPlanner + DistanceHeuristic (only with LIB_REF)."""

    def __init__(self, orc, radius, world=None):
        self.L = orc.L
        self.world = world
        self.h = self.L.ref_planner_create(radius, world.h if world else None)

    def close(self):
        if self.h:
            self.L.ref_planner_destroy(self.h)
            self.h = None

    def set_goal(self, end_pose, tol):
        self.L.ref_set_goal(self.h, _p(_f64(end_pose)), tol)

    def generate_grid(self, start, size, res, sample_orient=False, orient_res=360.0):
        self.L.ref_generate_grid(self.h, _p(_f64(start)), _p(_f64(size)), res, int(bool(sample_orient)), orient_res)

    def insert_nodes(self, poses):
        P = _f64(poses, 7)
        for i in range(len(P)):
            self.L.ref_insert_node(self.h, _p(P[i]))

    def insert_nodes_ids(self, poses):
        """-> id of the node insertNode returned for every pose (its own id when stored)"""
        P = _f64(poses, 7)
        return np.array([self.L.ref_insert_node_id(self.h, _p(P[i])) for i in range(len(P))], dtype=np.int64)

    def nodes(self):
        n = self.L.ref_num_nodes(self.h)
        P = np.zeros((n, 7))
        ids = np.zeros(n, dtype=np.int32)
        self.L.ref_get_nodes(self.h, _p(P), _p(ids))
        return P, ids

    def load_regular_grid(self, f1, f2, f3):
        self.L.ref_load_regular_grid(self.h, f1.encode(), f2.encode(), f3.encode())

    def save_search_space(self, fn):
        self.L.ref_save_search_space(self.h, fn.encode())

    def load_search_space(self, fn):
        self.L.ref_load_search_space(self.h, fn.encode())

    def connect_to_nn(self, hull1, hull2, orient_res):
        a = np.ascontiguousarray(hull1, dtype=np.float32).reshape(-1, 3)
        b = np.ascontiguousarray(hull2, dtype=np.float32).reshape(-1, 3)
        self.L.ref_connect_to_nn(self.h, _p(a), len(a), _p(b), len(b), orient_res)

    def maxnodes(self):
        return self.L.ref_maxnodes(self.h)

    def disconnect(self):
        self.L.ref_disconnect(self.h)

    def connect(self):
        return self.L.ref_connect(self.h)

    def collision_calls(self):
        return self.L.ref_collision_calls(self.h)

    def csr(self):
        n = self.L.ref_num_nodes(self.h)
        rp = np.zeros(n + 1, dtype=np.int64)
        E = self.L.ref_get_csr(self.h, _p(rp), None, 0)
        col = np.zeros(max(E, 1), dtype=np.int32)
        self.L.ref_get_csr(self.h, _p(rp), _p(col), E)
        return rp, col[:E]

    def astar(self, start_pose, cap=4096):
        path = np.zeros((cap, 10))
        stats = np.zeros(2, dtype=np.int64)
        k = self.L.ref_astar(self.h, _p(_f64(start_pose)), _p(path), cap, _p(stats))
        return path[:k], int(stats[0]), int(stats[1])


class RefCoveragePlanner(RefPlanner):
    """The reference's own PathPlanner driven through oracle/ref_harness.cpp's HarnessCoverage: the authored
    coverage semantics on the oracle, so that the reference's dynamic sampling loop can run (only with LIB_REF)."""

    def __init__(self, orc, radius, world, cam):
        self.L = orc.L
        self.world = world
        self.h = self.L.ref_coverage_planner_create(radius, world.h if world else None, _p(_f64(cam["mount_xyz"])),
                                                    _p(_f64(cam["mount_rpy"])), cam["hfov"], cam["vfov"], cam["near"], cam["far"])

    def set_model_points(self, pts):
        a = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        self.L.ref_coverage_set_model(self.h, _p(a), len(a))

    def set_clustering(self, tolerance, min_size, standoff):
        self.L.ref_coverage_set_clustering(self.h, tolerance, min_size, standoff)

    def dynamic_sampling(self, start, size, start_res, res_decrement):
        out = np.zeros(3, dtype=np.int64)
        self.L.ref_dynamic_sampling(self.h, _p(_f64(start)), _p(_f64(size)), start_res, res_decrement, _p(out))
        return int(out[0]), int(out[1]), int(out[2])


class Oracle:
    def __init__(self, path=LIB_ORACLE):
        if not os.path.exists(path):
            raise ImportError(f"{path} missing: run __graft_entry__.build()")
        L = ctypes.CDLL(path)
        self.L = L
        self.has_ref = hasattr(L, "ref_planner_create")
        L.orc_world_create.restype = vp
        L.orc_world_create.argtypes = [dbl, ctypes.c_int]
        L.orc_world_destroy.argtypes = [vp]
        L.orc_world_set_box.argtypes = [vp, vp, vp, ctypes.c_int]
        L.orc_world_set_keys.argtypes = [vp, vp, lng, ctypes.c_int]
        L.orc_world_prune.argtypes = [vp]
        L.orc_world_insert_pointcloud.argtypes = [vp, vp, lng, vp, dbl]
        L.orc_world_export_logodds.argtypes = [vp, vp, vp, vp]
        L.orc_world_num_tree_nodes.restype = lng
        L.orc_world_num_tree_nodes.argtypes = [vp]
        L.orc_world_build_dense.argtypes = [vp, vp, vp, vp, vp, ctypes.c_int]
        L.orc_world_export_boxes.restype = lng
        L.orc_world_export_boxes.argtypes = [vp, ctypes.c_int, vp, vp, lng]
        L.orc_world_export_dense.argtypes = [vp, vp, vp, vp, vp]
        L.orc_ray_lattice.argtypes = [vp, dbl, vp, vp, vp, vp]
        L.orc_ray_keys.restype = lng
        L.orc_ray_keys.argtypes = [vp, vp, vp, vp, lng]
        L.orc_line_status.argtypes = [vp, vp, vp, lng, vp]
        L.orc_line_status_bbox.argtypes = [vp, vp, vp, lng, vp, vp, vp, vp]
        L.orc_line_status_bbox_mt.argtypes = [vp, vp, vp, lng, vp, vp, ctypes.c_int]
        L.orc_cell_status_point.argtypes = [vp, vp, lng, vp]
        L.orc_cell_status_bbox.argtypes = [vp, vp, lng, vp, vp]
        L.orc_connect.restype = lng
        L.orc_connect.argtypes = [vp, vp, lng, dbl, vp, ctypes.c_int, vp, vp, lng, vp]
        L.orc_cast_rays.argtypes = [vp, vp, vp, lng, ctypes.c_int, vp, vp, vp, vp]
        L.orc_frustum_consts.argtypes = [dbl, dbl, dbl, dbl, vp]
        L.orc_frustum_planes.argtypes = [vp, vp, vp]
        L.orc_dedup_poses.restype = lng
        L.orc_dedup_poses.argtypes = [vp, lng, vp, vp]
        L.orc_visibility.restype = lng
        L.orc_visibility.argtypes = [vp, vp, lng, vp, vp, lng, vp, vp, lng, vp]
        if self.has_ref:
            L.ref_planner_create.restype = vp
            L.ref_planner_create.argtypes = [dbl, vp]
            L.ref_planner_destroy.argtypes = [vp]
            L.ref_set_goal.argtypes = [vp, vp, dbl]
            L.ref_generate_grid.argtypes = [vp, vp, vp, ctypes.c_float, ctypes.c_int, ctypes.c_float]
            L.ref_insert_node.argtypes = [vp, vp]
            L.ref_insert_node_id.restype = lng
            L.ref_insert_node_id.argtypes = [vp, vp]
            L.ref_num_nodes.restype = lng
            L.ref_num_nodes.argtypes = [vp]
            L.ref_get_nodes.argtypes = [vp, vp, vp]
            L.ref_disconnect.argtypes = [vp]
            L.ref_connect_to_nn.argtypes = [vp, vp, lng, vp, lng, ctypes.c_float]
            L.ref_maxnodes.restype = lng
            L.ref_maxnodes.argtypes = [vp]
            L.ref_load_regular_grid.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p]
            L.ref_save_search_space.argtypes = [vp, ctypes.c_char_p]
            L.ref_load_search_space.argtypes = [vp, ctypes.c_char_p]
            L.ref_connect.restype = lng
            L.ref_connect.argtypes = [vp]
            L.ref_collision_calls.restype = lng
            L.ref_collision_calls.argtypes = [vp]
            L.ref_get_csr.restype = lng
            L.ref_get_csr.argtypes = [vp, vp, vp, lng]
            L.ref_astar.restype = lng
            L.ref_astar.argtypes = [vp, vp, vp, lng, vp]
            if hasattr(L, "ref_coverage_planner_create"):
                L.ref_coverage_planner_create.restype = vp
                L.ref_coverage_planner_create.argtypes = [dbl, vp, vp, vp, dbl, dbl, dbl, dbl]
                L.ref_coverage_set_model.argtypes = [vp, vp, lng]
                L.ref_coverage_set_clustering.argtypes = [vp, dbl, ctypes.c_int, dbl]
                L.ref_dynamic_sampling.argtypes = [vp, vp, vp, dbl, dbl, vp]

    def world(self, resolution, treat_unknown=True):
        return World(self, resolution, treat_unknown)

    def planner(self, radius, world=None):
        assert self.has_ref
        return RefPlanner(self, radius, world)

    def coverage_planner(self, radius, world, cam):
        assert self.has_ref
        return RefCoveragePlanner(self, radius, world, cam)

    def ray_lattice(self, box, res):
        offs = [np.zeros(4096) for _ in range(3)]
        n = np.zeros(3, dtype=np.int32)
        self.L.orc_ray_lattice(_p(_f64(box)), res, _p(offs[0]), _p(offs[1]), _p(offs[2]), _p(n))
        return [offs[a][:n[a]].copy() for a in range(3)]

    def frustum_consts(self, hfov, vfov, near, far):
        out = np.zeros(4, dtype=np.float32)
        self.L.orc_frustum_consts(hfov, vfov, near, far, _p(out))
        return out

    def dedup_poses(self, poses):
        """Restated SearchSpace::insertNode over a batch -> (keep uint8[n], dup_of int64[n])."""
        P = _f64(poses, 7)
        keep = np.zeros(len(P), dtype=np.uint8)
        dup = np.zeros(len(P), dtype=np.int64)
        self.L.orc_dedup_poses(_p(P), len(P), _p(keep), _p(dup))
        return keep, dup

    def connect(self, world, xyz, radius, box=(0.5, 0.5, 0.3), mode=0):
        """Restated O(N^2) connectNodes -> (row_ptr, col_idx, counters[cand, same, steps, rays])."""
        a = _f64(xyz, 3)
        n = len(a)
        rp = np.zeros(n + 1, dtype=np.int64)
        cnt = np.zeros(4, dtype=np.int64)
        h = world.h if world is not None else None
        E = self.L.orc_connect(h, _p(a), n, radius, _p(_f64(box)), mode, _p(rp), None, 0, _p(cnt))
        col = np.zeros(max(E, 1), dtype=np.int32)
        self.L.orc_connect(h, _p(a), n, radius, _p(_f64(box)), mode, _p(rp), _p(col), E, _p(cnt))
        return rp, col[:E], cnt


class Probe:
    """CPU build of the engine's own dda_core.cuh (tests/cpu_probe)."""

    def __init__(self):
        L = ctypes.CDLL(LIB_PROBE)
        self.L = L
        L.probe_ray_lattice.restype = ctypes.c_int
        L.probe_ray_lattice.argtypes = [dbl, dbl, vp, ctypes.c_int]
        L.probe_line_status_bbox.argtypes = [dbl, vp, vp, vp, vp, ctypes.c_int, vp, vp, lng, vp, vp, vp]
        L.probe_cast_rays.argtypes = [dbl, vp, vp, vp, vp, vp, vp, lng, ctypes.c_int, vp, vp, vp]

    def ray_lattice(self, size, res):
        out = np.zeros(64)
        n = self.L.probe_ray_lattice(size, res, _p(out), 64)
        return out[:max(n, 0)].copy(), n

    def line_status_bbox(self, res, key_min, dims, occ, known, treat, starts, ends, box):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        out = np.zeros(len(s), dtype=np.uint8)
        steps = lng()
        self.L.probe_line_status_bbox(res, _p(km), _p(dm), _p(np.ascontiguousarray(occ)), _p(np.ascontiguousarray(known)),
                                      int(bool(treat)), _p(s), _p(e), len(s), _p(_f64(box)), _p(out), ctypes.byref(steps))
        return out, steps.value

    def cast_rays(self, res, key_min, dims, occ, known, origins, dirs, ignore_unknown=True, max_range=None):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        mr = None if max_range is None else _f64(max_range)
        hit = np.zeros(len(o), dtype=np.uint8)
        keys = np.zeros((len(o), 3), dtype=np.int32)
        self.L.probe_cast_rays(res, _p(km), _p(dm), _p(np.ascontiguousarray(occ)), _p(np.ascontiguousarray(known)), _p(o), _p(d),
                               len(o), int(bool(ignore_unknown)), _p(mr), _p(hit), _p(keys))
        return hit, keys
