"""ctypes binding of the host C++ mirror's test hooks (sspp_b200/host/src/capi.cpp)."""
import ctypes

import numpy as np

from sspp_b200.build import LIB_HOST

vp = ctypes.c_void_p
i64 = ctypes.c_int64


def _p(a):
    return None if a is None else a.ctypes.data_as(vp)


def _f64(a, cols=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a if cols is None else a.reshape(-1, cols)


class HostError(RuntimeError):
    pass


_L = None


def lib():
    global _L
    if _L is None:
        L = ctypes.CDLL(LIB_HOST)
        L.sspph_create.restype = vp
        L.sspph_create.argtypes = [ctypes.c_double, ctypes.c_int]
        L.sspph_destroy.argtypes = [vp]
        L.sspph_error.restype = ctypes.c_char_p
        L.sspph_error.argtypes = [vp]
        L.sspph_ok.argtypes = [vp]
        L.sspph_map_create.argtypes = [vp, ctypes.c_double, vp, vp, ctypes.c_int]
        L.sspph_map_set_box.argtypes = [vp, vp, vp, ctypes.c_int]
        L.sspph_map_insert_boxes.argtypes = [vp, vp, vp, vp, i64]
        L.sspph_map_upload_dense.argtypes = [vp, vp, vp]
        L.sspph_set_goal.argtypes = [vp, vp, ctypes.c_double]
        L.sspph_set_filter_by_validity.argtypes = [vp, ctypes.c_int]
        L.sspph_generate_grid.argtypes = [vp, vp, vp, ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_int]
        L.sspph_insert_nodes.argtypes = [vp, vp, i64]
        L.sspph_load_regular_grid.argtypes = [vp, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p]
        L.sspph_save_search_space.argtypes = [vp, ctypes.c_char_p]
        L.sspph_load_search_space.argtypes = [vp, ctypes.c_char_p]
        L.sspph_save_binary.argtypes = [vp, ctypes.c_char_p]
        L.sspph_load_binary.argtypes = [vp, ctypes.c_char_p]
        L.sspph_num_nodes.restype = i64
        L.sspph_num_nodes.argtypes = [vp]
        L.sspph_get_nodes.argtypes = [vp, vp, vp]
        L.sspph_remove_node.argtypes = [vp, vp]
        L.sspph_set_adjacency.argtypes = [vp, vp, vp, i64]
        L.sspph_disconnect.argtypes = [vp]
        L.sspph_connect.argtypes = [vp, vp, vp]
        L.sspph_get_csr.restype = i64
        L.sspph_get_csr.argtypes = [vp, vp, vp, i64]
        L.sspph_astar.restype = i64
        L.sspph_astar.argtypes = [vp, vp, vp, i64, vp, vp, i64]
        L.sspph_robot2sensor.argtypes = [vp, vp, vp, i64, vp]
        L.sspph_connect_to_nn.argtypes = [vp, vp, i64, vp, i64, ctypes.c_float]
        L.sspph_maxnodes.restype = i64
        L.sspph_maxnodes.argtypes = [vp]
        L.sspph_duplicate_edges.restype = i64
        L.sspph_duplicate_edges.argtypes = [vp]
        L.sspph_dynamic_sampling.argtypes = [vp, vp, vp, ctypes.c_double, ctypes.c_double, vp]
        L.sspph_set_clustering.argtypes = [vp, ctypes.c_double, ctypes.c_int, ctypes.c_double]
        L.sspph_cluster_labels.argtypes = [vp, vp, i64, vp]
        L.sspph_create_coverage.restype = vp
        L.sspph_create_coverage.argtypes = [ctypes.c_double, ctypes.c_int, vp, vp] + [ctypes.c_double] * 4
        L.sspph_set_model_points.argtypes = [vp, vp, i64]
        L.sspph_set_coverage_goal.argtypes = [vp, ctypes.c_double, ctypes.c_double]
        L.sspph_visible_from.restype = i64
        L.sspph_visible_from.argtypes = [vp, i64, vp, i64]
        L.sspph_get_sensor_poses.argtypes = [vp, vp]
        L.sspph_coverage_counters.argtypes = [vp, vp]
        _L = L
    return _L


class HostPlanner:
    """PathPlanner + DistanceHeuristic (+ GpuWorld when device >= 0) of the host mirror."""

    def __init__(self, radius, device=-1, coverage=None):
        """coverage: None, or dict(mount_xyz, mount_rpy, hfov, vfov, near, far) for a CoverageHeuristic planner"""
        self.L = lib()
        if coverage is None:
            self.h = self.L.sspph_create(radius, device)
        else:
            self.h = self.L.sspph_create_coverage(radius, device, _p(_f64(coverage["mount_xyz"])), _p(_f64(coverage["mount_rpy"])),
                                                  coverage["hfov"], coverage["vfov"], coverage["near"], coverage["far"])
        if not self.L.sspph_ok(self.h):
            msg = self.L.sspph_error(self.h).decode()
            self.L.sspph_destroy(self.h)
            self.h = None
            raise HostError(msg)

    def close(self):
        if self.h:
            self.L.sspph_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise HostError(self.L.sspph_error(self.h).decode())

    def map_create(self, res, key_min, dims, treat_unknown=True):
        km, dm = np.asarray(key_min, dtype=np.int32), np.asarray(dims, dtype=np.int32)
        self._ck(self.L.sspph_map_create(self.h, res, _p(km), _p(dm), int(bool(treat_unknown))))

    def map_set_box(self, pos, box, occupied):
        self._ck(self.L.sspph_map_set_box(self.h, _p(_f64(pos)), _p(_f64(box)), int(bool(occupied))))

    def map_insert_boxes(self, centres, sizes, occupied):
        c, s = _f64(centres, 3), _f64(sizes)
        o = np.ascontiguousarray(occupied, dtype=np.uint8)
        self._ck(self.L.sspph_map_insert_boxes(self.h, _p(c), _p(s), _p(o), len(s)))

    def map_upload_dense(self, occ, known):
        self._ck(self.L.sspph_map_upload_dense(self.h, _p(np.ascontiguousarray(occ, dtype=np.uint32)),
                                               _p(np.ascontiguousarray(known, dtype=np.uint32))))

    def set_goal(self, end_pose, tol):
        self.L.sspph_set_goal(self.h, _p(_f64(end_pose)), tol)

    def set_model_points(self, pts):
        a = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        self._ck(self.L.sspph_set_model_points(self.h, _p(a), len(a)))

    def set_coverage_goal(self, target_pct, tol_pct):
        self.L.sspph_set_coverage_goal(self.h, target_pct, tol_pct)

    def visible_from(self, index, cap=1 << 22):
        idx = np.zeros(cap, dtype=np.int32)
        k = self.L.sspph_visible_from(self.h, index, _p(idx), cap)
        if k < 0:
            raise HostError(self.L.sspph_error(self.h).decode())
        return idx[:k].copy()

    def sensor_poses(self):
        n = self.L.sspph_num_nodes(self.h)
        P = np.zeros((n, 7))
        self.L.sspph_get_sensor_poses(self.h, _p(P))
        return P

    def coverage_counters(self):
        c = np.zeros(2, dtype=np.int64)
        self.L.sspph_coverage_counters(self.h, _p(c))
        return int(c[0]), int(c[1])

    def connect_to_nn(self, hull1, hull2, orient_res):
        a = np.ascontiguousarray(hull1, dtype=np.float32).reshape(-1, 3)
        b = np.ascontiguousarray(hull2, dtype=np.float32).reshape(-1, 3)
        self._ck(self.L.sspph_connect_to_nn(self.h, _p(a), len(a), _p(b), len(b), orient_res))

    def maxnodes(self):
        return self.L.sspph_maxnodes(self.h)

    def duplicate_edges(self):
        return self.L.sspph_duplicate_edges(self.h)

    def dynamic_sampling(self, start, size, start_res, res_decrement):
        out = np.zeros(3, dtype=np.int64)
        self._ck(self.L.sspph_dynamic_sampling(self.h, _p(_f64(start)), _p(_f64(size)), start_res, res_decrement, _p(out)))
        return int(out[0]), int(out[1]), int(out[2])

    def set_clustering(self, tolerance, min_size, standoff):
        self._ck(self.L.sspph_set_clustering(self.h, tolerance, min_size, standoff))

    def cluster_labels(self, pts):
        a = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
        lab = np.zeros(len(a), dtype=np.int32)
        self._ck(self.L.sspph_cluster_labels(self.h, _p(a), len(a), _p(lab)))
        return lab

    def set_filter_by_validity(self, on):
        self.L.sspph_set_filter_by_validity(self.h, int(bool(on)))

    def generate_grid(self, start, size, res, sample_orient=False, orient_res=360.0, filtering=False):
        self._ck(self.L.sspph_generate_grid(self.h, _p(_f64(start)), _p(_f64(size)), res, int(bool(sample_orient)), orient_res,
                                            int(bool(filtering))))

    def insert_nodes(self, poses):
        P = _f64(poses, 7)
        self.L.sspph_insert_nodes(self.h, _p(P), len(P))

    def load_regular_grid(self, f1, f2, f3):
        self._ck(self.L.sspph_load_regular_grid(self.h, f1.encode(), f2.encode(), f3.encode()))

    def save_search_space(self, fn):
        self._ck(self.L.sspph_save_search_space(self.h, fn.encode()))

    def load_search_space(self, fn):
        self._ck(self.L.sspph_load_search_space(self.h, fn.encode()))

    def save_binary(self, fn):
        self._ck(self.L.sspph_save_binary(self.h, fn.encode()))

    def load_binary(self, fn):
        self._ck(self.L.sspph_load_binary(self.h, fn.encode()))

    def nodes(self):
        n = self.L.sspph_num_nodes(self.h)
        P = np.zeros((n, 7))
        ids = np.zeros(n, dtype=np.int32)
        self.L.sspph_get_nodes(self.h, _p(P), _p(ids))
        return P, ids

    def remove_node(self, pose):
        return bool(self.L.sspph_remove_node(self.h, _p(_f64(pose))))

    def set_adjacency(self, row_ptr, col):
        rp = np.ascontiguousarray(row_ptr, dtype=np.int64)
        cl = np.ascontiguousarray(col, dtype=np.int32)
        self._ck(self.L.sspph_set_adjacency(self.h, _p(rp), _p(cl), len(rp) - 1))

    def disconnect(self):
        self.L.sspph_disconnect(self.h)

    def connect(self):
        e, t = i64(), i64()
        self._ck(self.L.sspph_connect(self.h, ctypes.byref(e), ctypes.byref(t)))
        return e.value, t.value

    def csr(self):
        n = self.L.sspph_num_nodes(self.h)
        rp = np.zeros(n + 1, dtype=np.int64)
        E = self.L.sspph_get_csr(self.h, _p(rp), None, 0)
        col = np.zeros(max(E, 1), dtype=np.int32)
        self.L.sspph_get_csr(self.h, _p(rp), _p(col), E)
        return rp, col[:E]

    def astar(self, start_pose, cap=4096, cap_order=1 << 16):
        path = np.zeros((cap, 10))
        stats = np.zeros(2, dtype=np.int64)
        order = np.full(cap_order, -1, dtype=np.int32)
        k = self.L.sspph_astar(self.h, _p(_f64(start_pose)), _p(path), cap, _p(stats), _p(order), cap_order)
        if k < 0:
            raise HostError(self.L.sspph_error(self.h).decode())
        return path[:k], int(stats[0]), int(stats[1]), order[:int(stats[1])]


def robot2sensor(mount_xyz, mount_rpy, robot_poses):
    P = _f64(robot_poses, 7)
    out = np.zeros_like(P)
    lib().sspph_robot2sensor(_p(_f64(mount_xyz)), _p(_f64(mount_rpy)), _p(P), len(P), _p(out))
    return out
