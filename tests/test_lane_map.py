"""Host logic of the tile kernel's contraction (no GPU): the block-pair units and the lane -> (unit, row slice) map that
build_structure() hands to k_frame_normal, through the mcc_debug_lane_map test hook."""
import ctypes as C

import numpy as np
import pytest


def _lane_map(n_dist, rowstride):
    from multicam_calibration_b200 import capi
    L = capi.lib()
    lane_u = (C.c_byte * 32)()
    lane_q = (C.c_byte * 32)()
    partner = (C.c_byte * 24)()
    nun = C.c_int()
    L.mcc_debug_lane_map.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_byte), C.POINTER(C.c_byte), C.POINTER(C.c_byte), C.POINTER(C.c_int)]
    L.mcc_debug_lane_map.restype = C.c_int
    ksl = L.mcc_debug_lane_map(n_dist, rowstride, lane_u, lane_q, partner, C.byref(nun))
    assert ksl > 0
    return ksl, nun.value, np.array(lane_u[:], dtype=int), np.array(lane_q[:], dtype=int), np.array(partner[:], dtype=int).reshape(3, 8)


@pytest.mark.parametrize("n_dist", range(0, 9))
def test_units_cover_every_block_pair_once(built, n_dist):
    for rs in (15, 21):
        if rs == 15 and 4 + n_dist + 7 > 15:
            continue
        ksl, nun, lane_u, lane_q, partner = _lane_map(n_dist, rs)
        assert nun == (4 + n_dist + 7 + 2) // 3
        pairs = []
        for u in range(nun):
            pairs.append((u, u))
            for d in range(3):
                if partner[d, u] >= 0:
                    pairs.append(tuple(sorted((u, int(partner[d, u])))))
        assert len(pairs) == len(set(pairs)) == nun * (nun + 1) // 2, pairs
        # every (unit, slice) on exactly one lane; slice counts the reduction handles (2^n or 3 * 2^n)
        used = [(u, q) for u, q in zip(lane_u, lane_q) if u >= 0]
        assert sorted(used) == [(u, q) for u in range(nun) for q in range(ksl)]
        assert all((u >= 0) == (q >= 0) for u, q in zip(lane_u, lane_q))
        m = ksl
        while m % 2 == 0:
            m //= 2
        assert m in (1, 3) and nun * ksl <= 32


@pytest.mark.parametrize("n_dist,rs,ksl_expected", [(4, 15, 6), (8, 21, 4)])
def test_half_warps_are_bank_conflict_free(built, n_dist, rs, ksl_expected):
    """the two configurations the kernel has unrolled loops for: 5 units x 6 slices (15 columns) and 7 units x 4 slices
    (21 columns).  A 64-bit shared load is served per half-warp; within one, every lane must hit a different 8-byte
    bank (16 of them) for the unit's own block column and for each partner column."""
    ksl, nun, lane_u, lane_q, partner = _lane_map(n_dist, rs)
    assert ksl == ksl_expected
    for half in (range(0, 16), range(16, 32)):
        for role in range(4):
            banks = []
            for l in half:
                u, q = lane_u[l], lane_q[l]
                if u < 0:
                    continue
                b = u if role == 0 or partner[role - 1, u] < 0 else partner[role - 1, u]
                banks.append((rs * q + 3 * b) % 16)          # word address of column 0 of the block, row q; columns 1, 2 shift all alike
            assert len(banks) == len(set(banks)), (role, banks)
