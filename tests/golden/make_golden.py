"""Generates tests/golden/*.npz from the REFERENCE'S OWN planner sources (oracle/_ref/libsspp_ref.so,
built by oracle/Makefile from /root/reference) — run in the build container only:

    python tests/golden/make_golden.py

The reference cannot travel to the GPU box; these small vectors do.  Scenarios:
  dht_known_answer   src/distance_heuristic_test.cpp:52-117 exactly (grid 10x10x2 @1.0, 4 yaws, R 1.5,
                     no map, start (0,0,0), end (4,4,2), tolerance 1.0)
  reactive_sphere    src/reactive_planner_server.cpp:40-42,58-131 + launch/reactive_planner_test.launch:
                     octomap res 0.3, treat_unknown false, setFree 15x15x10 box, occupied sphere r=1.0 at
                     the origin (the launch file's point cloud is not in the repo), grid
                     10x10x0 @1.0 around grid_start = start (-2.12,1.2,0.2), R 1.5, goal (2,2,0) tol 0.5
  uav_fixture_R25    resources/SearchSpaceUAV_1.5m_1to4_NEW.txt first 2000 poses, R 2.5, no map
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402


def sphere_keys(res, radius, centre=(0.0, 0.0, 0.0)):
    """Voxels whose centre lies within `radius` of `centre` (stand-in for the missing sphere cloud)."""
    n = int(np.ceil(radius / res)) + 2
    k0 = [int(np.floor(c / res)) for c in centre]
    out = []
    for dz in range(-n, n + 1):
        for dy in range(-n, n + 1):
            for dx in range(-n, n + 1):
                k = (k0[0] + dx, k0[1] + dy, k0[2] + dz)
                c = [(k[a] + 0.5) * res for a in range(3)]
                if sum((c[a] - centre[a]) ** 2 for a in range(3)) <= radius * radius:
                    out.append([k[0] + 32768, k[1] + 32768, k[2] + 32768])
    return np.array(out, dtype=np.int32)


def main():
    O = oracle_lib.Oracle(oracle_lib.LIB_REF)

    # 1. distance_heuristic_test
    p = O.planner(1.5)
    p.set_goal([4, 4, 2, 0, 0, 0, 0], 1.0)
    p.generate_grid([0, 0, 0], [10, 10, 2], 1.0, True, 90.0)
    poses, ids = p.nodes()
    E = p.connect()
    rp, col = p.csr()
    path, created, expanded = p.astar([0, 0, 0, 0, 0, 0, 0])
    np.savez_compressed(os.path.join(HERE, "dht_known_answer.npz"), poses=poses, ids=ids, row_ptr=rp, col_idx=col,
                        path=path, created=created, expanded=expanded, connections=E)
    print("dht", len(poses), E, created, expanded, path[-1, 7])
    p.close()

    # 2. reactive sphere scene
    res = 0.3
    w = O.world(res, False)
    w.set_box([0, 0, 0], [15, 15, 10], False)
    keys = sphere_keys(res, 1.0)
    w.set_keys(keys, True)
    p = O.planner(1.5, w)
    p.set_goal([2, 2, 0, 0, 0, 0, 0], 0.5)
    p.generate_grid([-2.12, 1.2, 0.2], [10, 10, 0], 1.0, False, 90.0)
    poses, ids = p.nodes()
    E = p.connect()
    rp, col = p.csr()
    path, created, expanded = p.astar([-2.12, 1.2, 0.2, 0, 0, 0, 0])
    np.savez_compressed(os.path.join(HERE, "reactive_sphere.npz"), poses=poses, ids=ids, row_ptr=rp, col_idx=col, path=path,
                        created=created, expanded=expanded, connections=E, sphere_keys=keys,
                        collision_calls=p.collision_calls())
    print("reactive", len(poses), E, p.collision_calls(), created, expanded, len(path))
    p.close()
    w.close()

    # 3. UAV viewpoint fixture (first 2000 rows), no map
    src = "/root/reference/resources/SearchSpaceUAV_1.5m_1to4_NEW.txt"
    rows = np.loadtxt(src, max_rows=2000)
    p = O.planner(2.5)
    p.insert_nodes(rows)
    poses, ids = p.nodes()
    E = p.connect()
    rp, col = p.csr()
    # store a checksum of the adjacency instead of the full list (kept small)
    h = np.bitwise_xor.reduce((col.astype(np.uint64) + 1) * np.uint64(0x9E3779B97F4A7C15) ^
                              np.repeat(np.arange(len(rp) - 1, dtype=np.uint64), np.diff(rp)) * np.uint64(0xC2B2AE3D27D4EB4F))
    np.savez_compressed(os.path.join(HERE, "uav_fixture_R25.npz"), poses=poses, row_ptr=rp, col_hash=np.uint64(h),
                        connections=E, col_head=col[:4096])
    print("uav", len(poses), E, hex(int(h)))
    p.close()


if __name__ == "__main__" and "--extra" not in sys.argv:
    main()


def casarea_map():
    """resources/casarea_s.png through resources/map.yaml (res 0.05, origin (-16,-12,0), occupied when
    (255-gray)/255 > 0.65, free when < 0.196, else unknown; image row 0 is the top = largest y), extruded
    over z in [-1, 1) m: the S1m variant of SURVEY.md §8(d).  Returns key_min, dims, occ, known bit planes."""
    from PIL import Image
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from sspp_b200 import scenes
    a = np.array(Image.open("/root/reference/resources/casarea_s.png").convert("L")).astype(np.float64)
    p = (255.0 - a) / 255.0
    occ2 = (p > 0.65)[::-1]          # row 0 of the flipped image = smallest y
    free2 = (p < 0.196)[::-1]
    nz = 40
    occ = np.repeat(occ2[None], nz, axis=0)
    known = np.repeat((occ2 | free2)[None], nz, axis=0)
    key_min = np.array([32768 - 320, 32768 - 240, 32768 - 20], dtype=np.int32)   # origin (-16,-12,-1) at res 0.05
    dims = np.array([640, 480, nz], dtype=np.int32)
    return key_min, dims, scenes.pack_bits(occ), scenes.pack_bits(known), occ2, (occ2 | free2)


def make_casarea():
    O = oracle_lib.Oracle(oracle_lib.LIB_REF)
    key_min, dims, occ, known, occ2, known2 = casarea_map()
    w = O.world(0.05, True)
    w.build_dense(key_min, dims, occ, known)
    p = O.planner(1.5, w)
    p.set_goal([5, 5, 0.5, 0, 0, 0, 0], 0.6)
    p.generate_grid([0, 0, 0], [10, 10, 1], 1.0, False, 90.0)
    poses, ids = p.nodes()
    E = p.connect()
    rp, col = p.csr()
    path, created, expanded = p.astar([-2, -5, 0.5, 0, 0, 0, 0])
    np.savez_compressed(os.path.join(HERE, "casarea_s1m.npz"), occ2d=np.packbits(occ2, axis=1), known2d=np.packbits(known2, axis=1),
                        poses=poses, row_ptr=rp, col_idx=col, path=path, created=created, expanded=expanded, connections=E,
                        collision_calls=p.collision_calls())
    print("casarea", len(poses), E, p.collision_calls(), created, expanded, len(path))
    p.close()
    w.close()


def make_cam_fixture():
    rows = np.loadtxt("/root/reference/resources/SearchSpaceCam_1.5m_1to4_NEW.txt", max_rows=14765)
    sel = rows[::37][:400]   # spread over the whole file: 8 yaws, all heights
    np.savez_compressed(os.path.join(HERE, "cam_fixture_poses.npz"), poses=sel)
    print("cam poses", sel.shape)


if __name__ == "__main__" and "--extra" in sys.argv:
    make_casarea()
    make_cam_fixture()


def make_full_config2():
    """BASELINE configs[1] at its FULL size: all 14 765 row-aligned UAV / camera poses of
    resources/SearchSpaceUAV_1.5m_1to4_NEW.txt and SearchSpaceCam_1.5m_1to4_NEW.txt (read by
    src/pathplanner.cpp:204-269), and the graph the reference's own connectNodes builds on them (R 2.5, no map:
    the manager == NULL case).  The poses are stored as the files give them (three decimals), the adjacency as
    row_ptr + a checksum + its first 4096 columns.      python tests/golden/make_golden.py --full"""
    O = oracle_lib.Oracle(oracle_lib.LIB_REF)
    uav = np.loadtxt("/root/reference/resources/SearchSpaceUAV_1.5m_1to4_NEW.txt")
    cam = np.loadtxt("/root/reference/resources/SearchSpaceCam_1.5m_1to4_NEW.txt")
    assert uav.shape == cam.shape == (14765, 7)
    p = O.planner(2.5)
    p.insert_nodes(uav)
    poses, ids = p.nodes()
    E = p.connect()
    rp, col = p.csr()
    h = np.bitwise_xor.reduce((col.astype(np.uint64) + 1) * np.uint64(0x9E3779B97F4A7C15) ^
                              np.repeat(np.arange(len(rp) - 1, dtype=np.uint64), np.diff(rp)) * np.uint64(0xC2B2AE3D27D4EB4F))
    # list order is insertion order reversed: SearchSpaceNode::id counts insertNode calls (src/searchspace.cpp:64-67),
    # so ids[i] is the file row of list entry i — and of its camera pose
    np.savez_compressed(os.path.join(HERE, "config2_full.npz"), poses=poses, ids=ids, cam_rows=cam, row_ptr=rp.astype(np.int32),
                        col_hash=np.uint64(h), connections=E, col_head=col[:4096])
    assert np.array_equal(poses, uav[ids])
    print("config2 full", len(poses), "stored of", len(uav), "E", E, hex(int(h)))
    p.close()


if __name__ == "__main__" and "--full" in sys.argv:
    make_full_config2()
