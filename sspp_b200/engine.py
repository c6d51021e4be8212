"""ctypes binding of include/sspp_gpu.h.  Thin: every method is one C-ABI call.

This is a driver for tests/ and bench.py, not the product; the product is the shared library.  It
refuses to work without the built CUDA library — there is no CPU path behind it.
"""
import ctypes
import os

import numpy as np

from .build import LIB_GPU

c_i64 = ctypes.c_int64
c_vp = ctypes.c_void_p


class SsppError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sspp_gpu error {code}: {msg}")
        self.code = code


class ConnectParams(ctypes.Structure):
    _fields_ = [("radius", ctypes.c_double), ("box", ctypes.c_double * 3), ("use_map", ctypes.c_int32),
                ("position_mode", ctypes.c_int32), ("row_begin", c_i64), ("row_end", c_i64),
                ("use_pyramid", ctypes.c_int32), ("count_steps", ctypes.c_int32), ("count_only", ctypes.c_int32),
                ("reserved", ctypes.c_int32)]


class ConnectStats(ctypes.Structure):
    _fields_ = [("rows", c_i64), ("edges", c_i64), ("in_radius_pairs", c_i64), ("tested_pairs", c_i64),
                ("rays_per_pair", c_i64), ("pyramid_rows", c_i64), ("tile_rows", c_i64), ("box_pairs", c_i64),
                ("lattice_pairs", c_i64), ("voxel_steps", c_i64)]


class PeerBlocks(ctypes.Structure):
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("row_ptr_blocks", c_vp), ("col_blocks", c_vp),
                ("row_stride", c_i64), ("col_stride", c_i64)]


class FusedPeers(ctypes.Structure):
    _fields_ = [("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("slot", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("first_row", c_i64), ("slot_peers", c_vp), ("deg_peers", c_vp)]


class Frustum(ctypes.Structure):
    _fields_ = [("hfov_deg", ctypes.c_double), ("vfov_deg", ctypes.c_double), ("near_dist", ctypes.c_double),
                ("far_dist", ctypes.c_double), ("count_steps", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class SensorModel(ctypes.Structure):
    _fields_ = [("probability_hit", ctypes.c_double), ("probability_miss", ctypes.c_double), ("threshold_min", ctypes.c_double),
                ("threshold_max", ctypes.c_double), ("threshold_occupancy", ctypes.c_double)]


class VisibilityStats(ctypes.Structure):
    _fields_ = [("visible_pairs", c_i64), ("rays", c_i64), ("voxel_steps", c_i64)]


_lib = None


def load_library():
    """Loads sspp_b200/csrc/libsspp_gpu.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_GPU):
        raise ImportError(f"{LIB_GPU} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the engine has no CPU fallback)")
    L = ctypes.CDLL(LIB_GPU)
    sig = {
        "sspp_gpu_create": (ctypes.c_int, [ctypes.POINTER(c_vp), ctypes.c_int]),
        "sspp_gpu_destroy": (None, [c_vp]),
        "sspp_gpu_last_error": (ctypes.c_char_p, [c_vp]),
        "sspp_gpu_stream": (c_vp, [c_vp]),
        "sspp_gpu_synchronize": (ctypes.c_int, [c_vp]),
        "sspp_gpu_map_create": (ctypes.c_int, [c_vp, ctypes.c_double, c_vp, c_vp, ctypes.c_int]),
        "sspp_gpu_map_set_treat_unknown": (ctypes.c_int, [c_vp, ctypes.c_int]),
        "sspp_gpu_map_upload_boxes": (ctypes.c_int, [c_vp, c_vp, c_vp, c_vp, c_i64]),
        "sspp_gpu_map_set_box": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int]),
        "sspp_gpu_map_upload_dense": (ctypes.c_int, [c_vp, c_vp, c_vp]),
        "sspp_gpu_map_download_dense": (ctypes.c_int, [c_vp, c_vp, c_vp]),
        "sspp_gpu_map_set_sensor_model": (ctypes.c_int, [c_vp, ctypes.POINTER(SensorModel)]),
        "sspp_gpu_map_insert_pointcloud": (ctypes.c_int, [c_vp, c_vp, c_i64, c_vp, ctypes.c_double]),
        "sspp_gpu_map_download_logodds": (ctypes.c_int, [c_vp, c_vp]),
        "sspp_gpu_set_nodes": (ctypes.c_int, [c_vp, c_vp, c_i64]),
        "sspp_gpu_set_nodes_device": (ctypes.c_int, [c_vp, c_vp, c_i64]),
        "sspp_gpu_connect": (ctypes.c_int, [c_vp, ctypes.POINTER(ConnectParams), ctypes.POINTER(ConnectStats)]),
        "sspp_gpu_connect_fetch": (ctypes.c_int, [c_vp, c_vp, c_vp]),
        "sspp_gpu_connect_fetch_device": (ctypes.c_int, [c_vp, c_vp, c_vp]),
        "sspp_gpu_connect_row_classes": (ctypes.c_int, [c_vp, c_vp]),
        "sspp_gpu_balanced_cuts": (ctypes.c_int, [c_vp, ctypes.POINTER(ConnectParams), ctypes.c_int, ctypes.c_int, c_vp]),
        "sspp_gpu_assemble_csr_blocks": (ctypes.c_int, [c_vp, c_vp, c_vp, ctypes.c_int, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp]),
        "sspp_gpu_peer_buffer_create": (ctypes.c_int, [c_vp, c_i64, ctypes.POINTER(c_vp), c_vp]),
        "sspp_gpu_peer_buffer_open": (ctypes.c_int, [c_vp, c_vp, ctypes.POINTER(c_vp)]),
        "sspp_gpu_peer_buffer_release": (ctypes.c_int, [c_vp, c_vp]),
        "sspp_gpu_connect_push_csr": (ctypes.c_int, [c_vp, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
        "sspp_gpu_connect_to_peers": (ctypes.c_int, [c_vp, ctypes.POINTER(ConnectParams), ctypes.POINTER(PeerBlocks),
                                                     ctypes.POINTER(ConnectStats)]),
        "sspp_gpu_connect_result_device": (ctypes.c_int, [c_vp, ctypes.POINTER(c_vp), ctypes.POINTER(c_vp)]),
        "sspp_gpu_connect_fused": (ctypes.c_int, [c_vp, ctypes.POINTER(ConnectParams), ctypes.POINTER(FusedPeers),
                                                  ctypes.POINTER(ConnectStats)]),
        "sspp_gpu_pack_slots": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, ctypes.c_int32, c_vp, c_vp, c_i64, c_vp]),
        "sspp_gpu_connect_nodes_host": (ctypes.c_int, [c_vp, c_vp, c_i64, ctypes.POINTER(ConnectParams), c_vp, c_vp,
                                                       c_i64, ctypes.POINTER(ConnectStats)]),
        "sspp_gpu_connect_to_host": (ctypes.c_int, [c_vp, ctypes.POINTER(ConnectParams), c_vp, c_vp, c_i64,
                                                    ctypes.POINTER(ConnectStats)]),
        "sspp_gpu_connect_timings": (ctypes.c_int, [c_vp, c_vp]),
        "sspp_gpu_dedup_poses": (ctypes.c_int, [c_vp, c_vp, c_i64, ctypes.c_double, c_vp, c_vp, ctypes.POINTER(c_i64)]),
        "sspp_gpu_check_segments": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp]),
        "sspp_gpu_check_segments_noalloc": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, c_vp, c_vp]),
        "sspp_gpu_validity": (ctypes.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
        "sspp_gpu_cast_rays": (ctypes.c_int, [c_vp, c_vp, c_vp, c_i64, ctypes.c_int, c_vp, c_vp, c_vp]),
        "sspp_gpu_set_model_points": (ctypes.c_int, [c_vp, c_vp, c_i64]),
        "sspp_gpu_visibility": (ctypes.c_int, [c_vp, c_vp, c_i64, ctypes.POINTER(Frustum), ctypes.POINTER(VisibilityStats)]),
        "sspp_gpu_visibility_fetch": (ctypes.c_int, [c_vp, c_vp, c_vp]),
        "sspp_gpu_visibility_timings": (ctypes.c_int, [c_vp, c_vp]),
        "sspp_gpu_visibility_push": (ctypes.c_int, [c_vp, ctypes.c_int, ctypes.c_int, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
        "sspp_gpu_measure_l2_bandwidth": (ctypes.c_int, [c_vp, c_i64, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]),
        "sspp_gpu_launch_count": (c_i64, [c_vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(c_vp)


def _f64(a, cols=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if cols is not None:
        a = a.reshape(-1, cols)
    return a


class Engine:
    """One sspp_gpu context (one GPU, one stream)."""

    def __init__(self, device=0):
        self.L = load_library()
        h = c_vp()
        rc = self.L.sspp_gpu_create(ctypes.byref(h), device)
        if rc != 0:
            raise SsppError(rc, "sspp_gpu_create failed (no CUDA device?)")
        self.h = h
        self.user_dims = None

    def close(self):
        if getattr(self, "h", None):
            self.L.sspp_gpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise SsppError(rc, self.L.sspp_gpu_last_error(self.h).decode())

    # ---- map
    def map_create(self, resolution, key_min, dims, treat_unknown_as_occupied=True):
        km = np.asarray(key_min, dtype=np.int32)
        dm = np.asarray(dims, dtype=np.int32)
        self._ck(self.L.sspp_gpu_map_create(self.h, float(resolution), _p(km), _p(dm), int(bool(treat_unknown_as_occupied))))
        self.user_dims = tuple(int(d) for d in dm)

    def map_set_treat_unknown(self, flag):
        self._ck(self.L.sspp_gpu_map_set_treat_unknown(self.h, int(bool(flag))))

    def map_upload_boxes(self, centres, sizes, occupied):
        c = _f64(centres, 3)
        s = _f64(sizes)
        o = np.ascontiguousarray(occupied, dtype=np.uint8)
        self._ck(self.L.sspp_gpu_map_upload_boxes(self.h, _p(c), _p(s), _p(o), len(s)))

    def map_set_box(self, position, box, occupied):
        p = _f64(position)
        b = _f64(box)
        self._ck(self.L.sspp_gpu_map_set_box(self.h, _p(p), _p(b), int(bool(occupied))))

    def map_upload_dense(self, occ_bits, known_bits):
        o = np.ascontiguousarray(occ_bits, dtype=np.uint32)
        k = np.ascontiguousarray(known_bits, dtype=np.uint32)
        self._ck(self.L.sspp_gpu_map_upload_dense(self.h, _p(o), _p(k)))

    def map_download_dense(self):
        ux, uy, uz = self.user_dims
        wpr = (ux + 31) // 32
        o = np.zeros((uz, uy, wpr), dtype=np.uint32)
        k = np.zeros((uz, uy, wpr), dtype=np.uint32)
        self._ck(self.L.sspp_gpu_map_download_dense(self.h, _p(o), _p(k)))
        return o, k

    def map_set_sensor_model(self, hit=0.65, miss=0.4, tmin=0.12, tmax=0.97, tocc=0.7):
        m = SensorModel(hit, miss, tmin, tmax, tocc)
        self._ck(self.L.sspp_gpu_map_set_sensor_model(self.h, ctypes.byref(m)))

    def map_insert_pointcloud(self, points, origin, max_range=-1.0):
        a = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        self._ck(self.L.sspp_gpu_map_insert_pointcloud(self.h, _p(a), len(a), _p(_f64(origin)), float(max_range)))

    def map_download_logodds(self):
        ux, uy, uz = self.user_dims
        out = np.zeros((uz, uy, ux), dtype=np.float32)
        self._ck(self.L.sspp_gpu_map_download_logodds(self.h, _p(out)))
        return out

    # ---- nodes / connect
    def set_nodes(self, xyz):
        a = _f64(xyz, 3)
        self._ck(self.L.sspp_gpu_set_nodes(self.h, _p(a), len(a)))

    def set_nodes_device(self, dev_ptr, n):
        self._ck(self.L.sspp_gpu_set_nodes_device(self.h, c_vp(dev_ptr), n))

    @staticmethod
    def params(radius, box=(0.5, 0.5, 0.3), use_map=True, position_mode=0, rows=None, use_pyramid=True, count_steps=False):
        p = ConnectParams()
        p.radius = radius
        p.box[0], p.box[1], p.box[2] = box
        p.use_map = int(bool(use_map))
        p.position_mode = position_mode
        p.row_begin, p.row_end = (rows if rows is not None else (0, 0))
        p.use_pyramid = int(bool(use_pyramid))
        p.count_steps = int(bool(count_steps))
        return p

    def connect(self, **kw):
        p = self.params(**kw)
        s = ConnectStats()
        self._ck(self.L.sspp_gpu_connect(self.h, ctypes.byref(p), ctypes.byref(s)))
        return s

    def connect_fetch(self, stats, rp_out=None, col_out=None):
        """Last CSR to host; rp_out / col_out: caller's (e.g. pinned) arrays of at least rows+1 / edges entries."""
        rp = np.zeros(stats.rows + 1, dtype=np.int64) if rp_out is None else rp_out
        col = np.zeros(max(stats.edges, 1), dtype=np.int32) if col_out is None else col_out
        assert len(rp) >= stats.rows + 1 and len(col) >= stats.edges
        self._ck(self.L.sspp_gpu_connect_fetch(self.h, _p(rp), _p(col)))
        return rp[:stats.rows + 1], col[:stats.edges]

    def balanced_cuts(self, world, work_row_weight=8, **kw):
        """world + 1 cut points over the list indices with about equal work per block (set-up step, runs the count pass)."""
        p = self.params(**kw)
        cuts = np.zeros(world + 1, dtype=np.int64)
        self._ck(self.L.sspp_gpu_balanced_cuts(self.h, ctypes.byref(p), world, work_row_weight, _p(cuts)))
        return cuts

    def connect_row_classes(self, stats):
        """uint8 per row of the last connect: 0 no pair, 1 pyramid, 3 decided pair by pair (2 reserved)."""
        out = np.zeros(max(stats.rows, 1), dtype=np.uint8)
        self._ck(self.L.sspp_gpu_connect_row_classes(self.h, _p(out)))
        return out[:stats.rows]

    def connect_fetch_device(self, row_ptr_dev, col_idx_dev):
        """D2D copy of the last CSR into caller-owned device buffers (raw addresses)."""
        self._ck(self.L.sspp_gpu_connect_fetch_device(self.h, c_vp(row_ptr_dev), c_vp(col_idx_dev)))

    def assemble_csr_blocks(self, rp_blocks, col_blocks, world, row_stride, col_stride, row_base, rp_out, col_out, overflow):
        """all arguments are raw device addresses (int) except the three integers"""
        self._ck(self.L.sspp_gpu_assemble_csr_blocks(self.h, c_vp(rp_blocks), c_vp(col_blocks), world, row_stride, col_stride,
                                                     c_vp(row_base), c_vp(rp_out), c_vp(col_out), c_vp(overflow)))

    def peer_buffer_create(self, nbytes):
        """-> (device address, 64-byte IPC handle as bytes) of a buffer other processes on this box can open"""
        ptr, h = c_vp(), (ctypes.c_uint8 * 64)()
        self._ck(self.L.sspp_gpu_peer_buffer_create(self.h, int(nbytes), ctypes.byref(ptr), h))
        return ptr.value, bytes(h)

    def peer_buffer_open(self, handle):
        ptr, h = c_vp(), (ctypes.c_uint8 * 64).from_buffer_copy(handle)
        self._ck(self.L.sspp_gpu_peer_buffer_open(self.h, h, ctypes.byref(ptr)))
        return ptr.value

    def peer_buffer_release(self, ptr):
        self._ck(self.L.sspp_gpu_peer_buffer_release(self.h, c_vp(ptr)))

    def connect_push_csr(self, world, rank, edge_counts, row_base, rp_peers, col_peers, col_capacity, overflow):
        """edge_counts / row_base / overflow: raw device addresses; rp_peers / col_peers: lists of `world`
        device addresses (entry [rank] = this rank's own buffers)"""
        a = (c_vp * world)(*[c_vp(int(x)) for x in rp_peers])
        b = (c_vp * world)(*[c_vp(int(x)) for x in col_peers])
        self._ck(self.L.sspp_gpu_connect_push_csr(self.h, world, rank, c_vp(edge_counts), c_vp(row_base), a, b,
                                                  int(col_capacity), c_vp(overflow)))

    def connect_to_peers(self, world, rank, rp_blocks, col_blocks, row_stride, col_stride, **kw):
        """connect + chunked delivery of this rank's block into slot [rank] of every rank's block buffers
        (lists of `world` device addresses).  Returns the stats; raises SsppError (code 4) when the block does
        not fit col_stride — `needed_edges` on the exception says how many it has."""
        p = self.params(**kw)
        s = ConnectStats()
        a = (c_vp * world)(*[c_vp(int(x)) for x in rp_blocks])
        b = (c_vp * world)(*[c_vp(int(x)) for x in col_blocks])
        pb = PeerBlocks(world, rank, ctypes.cast(a, c_vp), ctypes.cast(b, c_vp), int(row_stride), int(col_stride))
        rc = self.L.sspp_gpu_connect_to_peers(self.h, ctypes.byref(p), ctypes.byref(pb), ctypes.byref(s))
        if rc:
            e = SsppError(rc, self.L.sspp_gpu_last_error(self.h).decode())
            e.needed_edges = int(s.edges)
            raise e
        return s

    def connect_fused(self, world, rank, slot, first_row, slot_peers, deg_peers, **kw):
        """this rank's rows with the exchange fused into the edge kernel: every finished row goes straight into the slot-format
        images of all ranks (lists of `world` device addresses).  Returns the stats (edges = this shard's); raises SsppError
        (code 4) when a row outgrows the slot."""
        p = self.params(**kw)
        s = ConnectStats()
        a = (c_vp * world)(*[c_vp(int(x)) for x in slot_peers])
        b = (c_vp * world)(*[c_vp(int(x)) for x in deg_peers])
        fp = FusedPeers(world, rank, int(slot), 0, int(first_row), ctypes.cast(a, c_vp), ctypes.cast(b, c_vp))
        self._ck(self.L.sspp_gpu_connect_fused(self.h, ctypes.byref(p), ctypes.byref(fp), ctypes.byref(s)))
        return s

    def pack_slots(self, slots, deg, n_rows, slot, row_ptr_out, col_out, col_capacity, overflow):
        """raw device addresses: a complete slot-format image -> whole-graph CSR on the engine's stream"""
        self._ck(self.L.sspp_gpu_pack_slots(self.h, c_vp(slots), c_vp(deg), int(n_rows), int(slot), c_vp(row_ptr_out), c_vp(col_out),
                                            int(col_capacity), c_vp(overflow)))

    def connect_result_device(self):
        a, b = c_vp(), c_vp()
        self._ck(self.L.sspp_gpu_connect_result_device(self.h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def connect_nodes_host(self, xyz, row_ptr, col_idx, **kw):
        """xyz (n,3) float64 host array, row_ptr/col_idx preallocated host arrays; returns stats."""
        p = self.params(**kw)
        s = ConnectStats()
        self._ck(self.L.sspp_gpu_connect_nodes_host(self.h, _p(xyz), len(xyz), ctypes.byref(p), _p(row_ptr), _p(col_idx),
                                                   len(col_idx), ctypes.byref(s)))
        return s

    def connect_to_host(self, row_ptr, col_idx, **kw):
        """connect on the nodes already set, CSR into preallocated host arrays; returns stats."""
        p = self.params(**kw)
        s = ConnectStats()
        self._ck(self.L.sspp_gpu_connect_to_host(self.h, ctypes.byref(p), _p(row_ptr), _p(col_idx), len(col_idx), ctypes.byref(s)))
        return s

    def connect_timings(self):
        t = np.zeros(8, dtype=np.float32)
        self._ck(self.L.sspp_gpu_connect_timings(self.h, _p(t)))
        return t

    # ---- node store
    def dedup_poses(self, poses, tolerance=1e-12, want_dup_of=True):
        """Batched SearchSpace::insertNode: (n,7) poses in insertion order -> (keep uint8[n], dup_of int64[n] or None, stored)."""
        P = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 7)
        keep = np.zeros(len(P), dtype=np.uint8)
        dup = np.zeros(len(P), dtype=np.int64) if want_dup_of else None
        stored = c_i64(0)
        self._ck(self.L.sspp_gpu_dedup_poses(self.h, _p(P), len(P), float(tolerance), _p(keep),
                                             _p(dup) if dup is not None else None, ctypes.byref(stored)))
        return keep, dup, stored.value

    # ---- queries
    def check_segments(self, starts, ends, box=(0.5, 0.5, 0.3)):
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        b = _f64(box)
        out = np.zeros(len(s), dtype=np.uint8)
        self._ck(self.L.sspp_gpu_check_segments(self.h, _p(s), _p(e), len(s), _p(b), _p(out)))
        return out

    def check_segments_noalloc(self, starts, ends, box=(0.5, 0.5, 0.3)):
        """Same query through the entry point that allocates no device memory (at most
        SSPP_NOALLOC_MAX_SEGMENTS = 16384 segments per call)."""
        s = _f64(starts, 3)
        e = _f64(ends, 3)
        b = _f64(box)
        out = np.zeros(len(s), dtype=np.uint8)
        self._ck(self.L.sspp_gpu_check_segments_noalloc(self.h, _p(s), _p(e), len(s), _p(b), _p(out)))
        return out

    def validity(self, points, box=(0.5, 0.5, 0.3)):
        a = _f64(points, 3)
        b = _f64(box)
        out = np.zeros(len(a), dtype=np.uint8)
        self._ck(self.L.sspp_gpu_validity(self.h, _p(a), len(a), _p(b), _p(out)))
        return out

    def cast_rays(self, origins, dirs, ignore_unknown=True, max_range=None):
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        mr = None if max_range is None else _f64(max_range)
        hit = np.zeros(len(o), dtype=np.uint8)
        keys = np.zeros((len(o), 3), dtype=np.int32)
        self._ck(self.L.sspp_gpu_cast_rays(self.h, _p(o), _p(d), len(o), int(bool(ignore_unknown)), _p(mr), _p(hit), _p(keys)))
        return hit, keys

    # ---- visibility
    def set_model_points(self, xyz):
        a = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        self._ck(self.L.sspp_gpu_set_model_points(self.h, _p(a), len(a)))

    def visibility(self, poses, hfov_deg, vfov_deg, near, far, fetch=True, count_steps=False, vp_out=None, idx_out=None):
        """vp_out / idx_out: caller's (e.g. pinned) arrays for the result; idx_out must hold the visible pairs."""
        P = poses if (isinstance(poses, np.ndarray) and poses.dtype == np.float64 and poses.flags.c_contiguous) else _f64(poses, 7)
        fr = Frustum(hfov_deg, vfov_deg, near, far, int(bool(count_steps)), 0)
        st = VisibilityStats()
        nv = P.size // 7
        self._ck(self.L.sspp_gpu_visibility(self.h, _p(P), nv, ctypes.byref(fr), ctypes.byref(st)))
        if not fetch:
            return st
        vp = np.zeros(nv + 1, dtype=np.int64) if vp_out is None else vp_out
        idx = np.zeros(max(st.visible_pairs, 1), dtype=np.int32) if idx_out is None else idx_out
        assert len(vp) >= nv + 1 and len(idx) >= st.visible_pairs
        self._ck(self.L.sspp_gpu_visibility_fetch(self.h, _p(vp), _p(idx)))
        return st, vp[:nv + 1], idx[:st.visible_pairs]

    def visibility_push(self, world, rank, pair_counts, view_base, vp_peers, idx_peers, idx_capacity, overflow):
        """as connect_push_csr, for the last visibility result (raw device addresses; lists of `world` peer buffers)"""
        a = (c_vp * world)(*[c_vp(int(x)) for x in vp_peers])
        b = (c_vp * world)(*[c_vp(int(x)) for x in idx_peers])
        self._ck(self.L.sspp_gpu_visibility_push(self.h, world, rank, c_vp(pair_counts), c_vp(view_base), a, b,
                                                 int(idx_capacity), c_vp(overflow)))

    def visibility_timings(self):
        t = np.zeros(8, dtype=np.float32)
        self._ck(self.L.sspp_gpu_visibility_timings(self.h, _p(t)))
        return t

    # ---- measurement
    def stream(self):
        return self.L.sspp_gpu_stream(self.h)

    def synchronize(self):
        self._ck(self.L.sspp_gpu_synchronize(self.h))

    def l2_bandwidth(self, nbytes=32 << 20, iters=20):
        g = ctypes.c_double()
        self._ck(self.L.sspp_gpu_measure_l2_bandwidth(self.h, nbytes, iters, ctypes.byref(g)))
        return g.value

    def launch_count(self):
        return int(self.L.sspp_gpu_launch_count(self.h))
