"""Integer arithmetic of the edge kernel's walk stage (sspp_b200/csrc/connect_kernels.cuh, "stage 3"; host side
ray_order_stride in sspp_b200/csrc/sspp_gpu.cu), restated in Python and checked exhaustively over its whole domain.

The kernel replaces three divisions by multiplications:
  lane / P            = (lane * (65536 / P + 1)) >> 16           for lane < 32, 1 <= P <= 32   (c_turn.magic)
  (t * stride) % R    = x - R * umulhi(x, 2^32 / R + 1), x = t * stride, t < R <= 1024           (ConnectArgs::ray_magic)
and walks the rays of a pair in the order t -> (t * stride) % R, which must visit every ray exactly once.
A wrong constant here would not crash: it would silently skip or repeat rays, so the domain is small enough to enumerate.
"""
from math import gcd


def ray_order_stride(rays):
    """Mirror of ray_order_stride() in sspp_gpu.cu (without the SSPP_RAY_STRIDE override)."""
    if rays < 3 or rays > 1024:
        return 1
    s = max(2, (rays * 3 + 9) // 10)
    while gcd(s, rays) != 1:
        s += 1
    return s if s < rays else 1


def test_lane_to_item_magic():
    for p in range(1, 33):
        magic = 65536 // p + 1
        rays_per_turn = 32 // p
        assert 1 <= rays_per_turn and p * rays_per_turn <= 32
        for lane in range(32):
            assert (lane * magic) >> 16 == lane // p


def test_ray_order_is_a_permutation_and_the_magic_modulo_is_exact():
    for r in range(1, 1025):
        s = ray_order_stride(r)
        assert 1 <= s and (s == 1 or (s < r and gcd(s, r) == 1))
        if s == 1:
            continue
        magic = 0xFFFFFFFF // r + 1
        seen = set()
        for t in range(r):
            x = t * s
            assert x < (1 << 32) // r  # the range in which umulhi(x, magic) == x // r
            q = x - r * ((x * magic) >> 32)
            assert q == x % r
            seen.add(q)
        assert len(seen) == r


def test_first_turn_samples_the_whole_lattice():
    """27 rays = 3 x 3 x 3 offsets, ray q = (ix * 3 + iy) * 3 + iz: the first nine rays of the strided order touch every
    offset value of every axis (the lattice order would stay on the ix = 0 face)."""
    r, s = 27, ray_order_stride(27)
    first = [(t * s) % r for t in range(9)]
    for axis in range(3):
        assert {(q // 3 ** (2 - axis)) % 3 for q in first} == {0, 1, 2}


def _lock_step(rays, pair_lanes, blocked_items):
    """Python model of the turn loop: returns (rays walked per pair, pairs found blocked, turns)."""
    s = ray_order_stride(rays)
    open_ = sorted(pair_lanes)
    walked = {k: [] for k in open_}
    blocked = set()
    q0 = turns = 0
    while open_:
        p = len(open_)
        k = min(rays - q0, 32 // p)
        hit = set()
        for lane in range(p * k):
            t = lane // p
            kk = open_[lane - t * p]
            q = ((q0 + t) * s) % rays
            walked[kk].append(q)
            if (kk, q) in blocked_items:
                hit.add(kk)
        turns += 1
        blocked |= hit
        open_ = [x for x in open_ if x not in hit]
        q0 += k
        if q0 >= rays:
            break
    return walked, blocked, turns


def test_lock_step_turns_walk_every_ray_of_a_free_pair_once_and_stop_blocked_pairs():
    import random
    rnd = random.Random(7)
    for _ in range(3000):
        rays = rnd.choice([2, 3, 8, 12, 27, 48, 100, 245])
        lanes = rnd.sample(range(32), rnd.randint(1, 32))
        density = rnd.choice([0.0, 0.02, 0.3])
        items = {(k, q) for k in lanes for q in range(rays) if rnd.random() < density}
        walked, blocked, turns = _lock_step(rays, lanes, items)
        for k in lanes:
            assert len(walked[k]) == len(set(walked[k]))  # no ray twice
            truth = any((k, q) in items for q in range(rays))
            assert (k in blocked) == truth  # the verdict is "any ray not free", whatever the order
            if not truth:
                assert sorted(walked[k]) == list(range(rays))
        assert turns <= rays
