"""CPU tests of the segmented-envelope planner (csrc/nd_plan.cuh) through its host entry point m3r_debug_nd_plan: the
schedule it emits for ring / chain / loop-closure keyframe graphs is checked against an independent symbolic Cholesky
of the block structure in the plan's elimination order (plain Python sets)."""
import ctypes as C

import numpy as np
import pytest

from mast3r_slam_ysh_b200 import _lib, synthetic as syn

BLK, POSE, MAXBLK, SCHED, MAXSEG, MAXBORDER, MAXLANES = 32, 7, 112, 8, 8, 4, 6
# NdPlan offsets (nd_plan.cuh)
HDR = 0
STAGE_BEGIN = 8
ORDER = STAGE_BEGIN + MAXSEG + 2
POS = ORDER + MAXBLK
STAGE_OF = POS + MAXBLK
SCHEDO = STAGE_OF + MAXBLK
LANE = SCHEDO + MAXBLK * SCHED
ADJS = LANE + (MAXSEG + 1) * MAXBLK
LROW = ADJS + 4 * MAXBLK
INS = LROW + 4 * MAXBLK
EXP = INS + 4
TOTAL = EXP + MAXSEG * (1 + MAXBORDER)


def block_adj(num_kf, edges):
    """Tile structure of the assembled system as index_prep_kernel marks it (gn_solve.cu): pose a's rows 7a..7a+6."""
    n = POSE * (num_kf - 1)
    nblk = (n + BLK - 1) // BLK
    adj = np.zeros((nblk, 4), dtype=np.uint32)

    def mark(ra, rc):
        hi, lo = max(ra, rc), min(ra, rc)
        adj[hi, lo >> 5] |= np.uint32(1 << (lo & 31))

    def blocks(a):
        return (POSE * a) // BLK, (POSE * a + POSE - 1) // BLK

    for a in range(num_kf - 1):
        r0, r1 = blocks(a)
        mark(r0, r0); mark(r1, r0); mark(r1, r1)
    for i, j in edges:
        a, o = max(i, j) - 1, min(i, j) - 1
        if o < 0:
            continue
        for ra in blocks(a):
            for ro in blocks(o):
                mark(ra, ro)
    return adj.view(np.int32), nblk


def _pairs(edges):  # synthetic.*_edges return (ii, jj)
    return [(int(i), int(j)) for i, j in zip(*edges)]


def run_plan(adj, nblk, P=0, sw=0):
    lib = _lib.load()
    assert lib.m3r_debug_nd_plan_ints() == TOTAL
    out = np.zeros(TOTAL, dtype=np.int32)
    adj = np.ascontiguousarray(adj)
    segs = lib.m3r_debug_nd_plan(adj.ctypes.data_as(C.c_void_p), nblk, P, sw, out.ctypes.data_as(C.c_void_p))
    return segs, out


def bits(words, nblk):
    return {b for b in range(nblk) if (int(words[b >> 5]) & 0xffffffff) >> (b & 31) & 1}


def check_plan(adj, nblk, segs, plan):
    adj = adj.view(np.uint32)
    assert plan[HDR] == segs and 2 <= segs <= MAXSEG
    order = [int(x) for x in plan[ORDER:ORDER + nblk]]
    assert sorted(order) == list(range(nblk))
    stage_of = plan[STAGE_OF:STAGE_OF + nblk]
    sb = plan[STAGE_BEGIN:STAGE_BEGIN + segs + 2]
    assert sb[0] == 0 and sb[segs + 1] == nblk
    for s in range(segs + 1):
        assert all(stage_of[k] == s for k in order[sb[s]:sb[s + 1]])
    in_s = bits(plan[INS:INS + 4], nblk)
    assert in_s == {k for k in range(nblk) if stage_of[k] == segs}
    assert order[-1] == nblk - 1  # the (possibly partial) last block is the last pivot
    # independent symbolic elimination in the plan's order
    struct = {r: {r} for r in range(nblk)}
    for r in range(nblk):
        for c in bits(adj[r], nblk):
            struct[r].add(c); struct[c].add(r)
    pos = {k: i for i, k in enumerate(order)}
    lrow_true = {r: set() for r in range(nblk)}
    for s in range(segs + 1):
        active = set()
        lanes = {}
        for k in order[sb[s]:sb[s + 1]]:
            members = {r for r in struct[k] if pos[r] >= pos[k]}
            for r in members - {k}:
                struct[r] |= members
                lrow_true[r].add(k)
            sc = plan[SCHEDO + k * SCHED: SCHEDO + (k + 1) * SCHED]
            m = int(sc[0])
            listed = [int(x) for x in sc[2:2 + m]]
            assert listed[0] == k and m <= MAXLANES
            assert [pos[r] for r in listed] == sorted(pos[r] for r in listed)  # elimination order
            assert members <= set(listed), (s, k, members, listed)
            new = {listed[i] for i in range(m) if (int(sc[1]) >> i) & 1}
            assert new == set(listed) - active
            active |= set(listed)
            for r in listed:  # lanes: one per active block, no two active blocks share one
                lanes[r] = int(plan[LANE + s * MAXBLK + r])
            assert len({lanes[r] for r in active}) == len(active) and max(lanes[r] for r in active) < MAXLANES
            # an interior only touches itself and separators
            assert all(stage_of[r] in (s, segs) for r in listed)
            active.discard(k)
        if s < segs:
            nb = int(plan[EXP + s * (1 + MAXBORDER)])
            border = [int(x) for x in plan[EXP + s * (1 + MAXBORDER) + 1: EXP + s * (1 + MAXBORDER) + 1 + nb]]
            assert set(border) == active and border == sorted(border) and set(border) <= in_s
    for r in range(nblk):
        assert bits(plan[LROW + 4 * r: LROW + 4 * r + 4], nblk) == lrow_true[r], r
    # separator stage loads every separator tile that is assembled or filled
    for a in in_s:
        want = {b for b in in_s if b <= a and (b in struct[a])}
        assert bits(plan[ADJS + 4 * a: ADJS + 4 * a + 4], nblk) >= want - {a} | ({a} & want)
    return max(sb[s + 1] - sb[s] for s in range(segs)) + (sb[segs + 1] - sb[segs])


@pytest.mark.parametrize("nkf", [64, 128, 256, 512])
def test_ring_graphs_split_and_schedule_is_a_valid_elimination(nkf):
    edges = _pairs(syn.ring_edges(nkf, 8))
    adj, nblk = block_adj(nkf, edges)
    segs, plan = run_plan(adj, nblk)
    assert segs >= 2, "a ring with window 8 must split"
    chain = check_plan(adj, nblk, segs, plan)
    assert chain < 0.75 * nblk, (chain, nblk)


def test_every_valid_candidate_is_a_valid_elimination():
    edges = _pairs(syn.ring_edges(128, 8))
    adj, nblk = block_adj(128, edges)
    valid = 0
    for P in range(2, 9):
        for sw in (1, 2, 3):
            segs, plan = run_plan(adj, nblk, P, sw)
            if segs:
                assert segs == P
                check_plan(adj, nblk, segs, plan)
                valid += 1
            if sw == 1:
                assert segs == 0  # one block (32 rows) cannot separate a window of 8 keyframes (up to 62 rows apart)
    assert valid >= 6


def test_open_chain_splits_and_dense_or_random_graphs_do_not():
    chain = [(i, j) for i in range(200) for j in range(i + 1, min(i + 5, 200))]
    adj, nblk = block_adj(200, chain)
    segs, plan = run_plan(adj, nblk)
    assert segs >= 2
    check_plan(adj, nblk, segs, plan)
    lc = _pairs(syn.loop_closure_edges(256))
    adj, nblk = block_adj(256, lc)
    assert run_plan(adj, nblk)[0] == 0
    dense = [(i, j) for i in range(40) for j in range(i + 1, 40)]
    adj, nblk = block_adj(40, dense)
    assert run_plan(adj, nblk)[0] == 0


def test_random_chains_with_extra_edges_never_yield_an_invalid_plan():
    """40 random graphs: a chain or ring with a random window plus a few random long-range edges.  Whatever the planner
    decides, a plan it returns must be a valid elimination (and most plain chains must split)."""
    rng = np.random.default_rng(7)
    split = 0
    for trial in range(40):
        nkf = int(rng.integers(45, 300))
        win = int(rng.integers(1, 10))
        ring = bool(rng.integers(0, 2))
        edges = [(i, (i + k) % nkf) for i in range(nkf) for k in range(1, win + 1) if ring or i + k < nkf]
        for _ in range(int(rng.integers(0, 4))):  # loop closures
            a, b = (int(x) for x in rng.integers(0, nkf, 2))
            if a != b:
                edges.append((a, b))
        adj, nblk = block_adj(nkf, edges)
        segs, plan = run_plan(adj, nblk)
        if segs:
            check_plan(adj, nblk, segs, plan)
            split += 1
    assert split >= 10

