"""Broadcast / reduce plans of the drivers on the process grids the benchmarks use (2x2, 2x4,
4x2, 1x4), checked without any communication: the plans are pure index arithmetic
(BaseMatrix::bcastPlan / reducePlan through the list builders the drivers call), so one
process can ask for them and replay every rank's tile inventory.  For every step of every
driver: each rank ends up holding every tile its local update reads, the root of a broadcast
is the tile's owner, nobody sends to itself, and a reduction is rooted at the owner of the C
tile with exactly the ranks holding partial sums as sources.  CPU only (no GPU, no NCCL).
"""
import pytest

GRIDS = [(2, 2), (2, 4), (4, 2), (1, 4)]


@pytest.fixture(scope="module")
def sl():
    from slate_d35_b200 import host, lib
    if lib().te_device_count() > 0:
        # with a GPU present a multi-rank initialize() wants a NCCL id; this test is about the
        # host-side planning only and runs where the "not gpu" suite runs
        pytest.skip("host-only planning test")
    return host


def owner(M, sl, i, j):
    return sl.tile_rank(M, i, j)


def replay(ops, which, have, world, M, sl):
    """apply the broadcasts of list `which`: have[r] gains tile (i, j) for every destination r"""
    for w, i, j, root, dests in ops:
        if w != which:
            continue
        assert root == owner(M, sl, i, j)
        assert root not in dests and len(set(dests)) == len(dests)
        assert all(0 <= d < world for d in dests)
        assert (i, j) in have[root]
        for d in dests:
            have[d].add((i, j))


def local_tiles(M, sl, mt, nt, world, lower=False):
    have = [set() for _ in range(world)]
    for i in range(mt):
        for j in range(nt):
            if lower and j > i:
                continue
            have[owner(M, sl, i, j)].add((i, j))
    return have


@pytest.mark.parametrize("p,q", GRIDS)
def test_gemm_and_gemmA_plans(sl, p, q):
    world = p * q
    sl.initialize(0, 0, world, None)
    nb, mt, nt, kt = 8, 7, 6, 5
    A = sl.Matrix(mt * nb - 3, kt * nb - 2, nb, p, q, origin="N")
    B = sl.Matrix(kt * nb - 2, nt * nb - 1, nb, p, q, origin="N")
    Cm = sl.Matrix(mt * nb - 3, nt * nb - 1, nb, p, q, origin="N")
    for k in range(kt):
        ops = sl.plan_gemm_step(A, B, Cm, k)
        ha, hb = local_tiles(A, sl, mt, kt, world), local_tiles(B, sl, kt, nt, world)
        replay(ops, 0, ha, world, A, sl)
        replay(ops, 1, hb, world, B, sl)
        for i in range(mt):
            for j in range(nt):
                r = owner(Cm, sl, i, j)
                assert (i, k) in ha[r] and (k, j) in hb[r]
    # gemmA: B(:, kc) to the owners of the block columns of A, C(:, kc) reduced onto its owners
    for kc in range(nt):
        ops = sl.plan_gemmA_step(A, B, Cm, kc)
        hb = local_tiles(B, sl, kt, nt, world)
        replay(ops, 0, hb, world, B, sl)
        for i in range(mt):
            for j in range(kt):
                assert (j, kc) in hb[owner(A, sl, i, j)]
        reduces = {(i, j): (root, src) for w, i, j, root, src in ops if w == 2}
        for i in range(mt):
            holders = {owner(A, sl, i, j) for j in range(kt)}
            root = owner(Cm, sl, i, kc)
            if holders - {root}:
                assert reduces[(i, kc)] == (root, sorted(holders - {root}))
            else:
                assert (i, kc) not in reduces
    sl.finalize()


@pytest.mark.parametrize("p,q", GRIDS)
def test_potrf_hemm_getrf_plans(sl, p, q):
    world = p * q
    sl.initialize(0, 0, world, None)
    nb, nt = 8, 9
    n = nt * nb - 3
    H = sl.HermitianMatrix("L", n, nb, p, q, origin="N")
    for k in range(nt):
        ops = sl.plan_potrf_step(H, k)
        have = local_tiles(H, sl, nt, nt, world, lower=True)
        replay(ops, 0, have, world, H, sl)
        for i in range(k + 1, nt):                       # trsm of A(i, k) needs A(k, k)
            assert (k, k) in have[owner(H, sl, i, k)]
        replay(ops, 1, have, world, H, sl)
        for j in range(k + 1, nt):                       # A(i, j) -= A(i, k) A(j, k)^H
            for i in range(j, nt):
                r = owner(H, sl, i, j)
                assert (i, k) in have[r] and (j, k) in have[r]

    # hemm, side Left, lower storage: C(i, :) += A(i, k) B(k, :)
    nrt = 4
    B = sl.Matrix(n, nrt * nb - 1, nb, p, q, origin="N")
    Cm = sl.Matrix(n, nrt * nb - 1, nb, p, q, origin="N")
    for k in range(nt):
        ops = sl.plan_hemm_step(H, B, Cm, k)
        ha, hb = local_tiles(H, sl, nt, nt, world, lower=True), local_tiles(B, sl, nt, nrt, world)
        replay(ops, 0, ha, world, H, sl)
        replay(ops, 1, hb, world, B, sl)
        for i in range(nt):
            need = (i, k) if i >= k else (k, i)
            for j in range(nrt):
                r = owner(Cm, sl, i, j)
                assert need in ha[r] and (k, j) in hb[r]

    # getrf_nopiv on a rectangular matrix
    mt, ntg = 8, 6
    G = sl.Matrix(mt * nb - 2, ntg * nb, nb, p, q, origin="N")
    for k in range(min(mt, ntg)):
        ops = sl.plan_getrf_nopiv_step(G, k)
        have = local_tiles(G, sl, mt, ntg, world)
        replay(ops, 0, have, world, G, sl)
        for i in range(k + 1, mt):
            assert (k, k) in have[owner(G, sl, i, k)]
        for j in range(k + 1, ntg):
            assert (k, k) in have[owner(G, sl, k, j)]
        replay(ops, 1, have, world, G, sl)
        replay(ops, 2, have, world, G, sl)
        for i in range(k + 1, mt):
            for j in range(k + 1, ntg):
                r = owner(G, sl, i, j)
                assert (i, k) in have[r] and (k, j) in have[r]
    sl.finalize()


@pytest.mark.parametrize("p,q", GRIDS)
def test_pair_messages_match_on_both_sides(sl, p, q):
    """BaseMatrix::bcastPairs (one message per root / destination pair, the opt-in coalesced
    broadcast): what r sends to d is exactly what d expects from r -- same tiles, same order,
    same byte count -- and together the messages are the plan, nothing more."""
    world = p * q
    sl.initialize(0, 0, world, None)
    nb, nt = 8, 9
    n = nt * nb - 3
    H = sl.HermitianMatrix("L", n, nb, p, q, origin="N")
    A = sl.Matrix(n, n, nb, p, q, origin="N")
    B = sl.Matrix(n, n, nb, p, q, origin="N")
    Cm = sl.Matrix(n, n, nb, p, q, origin="N")

    def check(kind, mats, k, which, plan_ops):
        msgs = [sl.plan_pairs(kind, mats, k, which, r) for r in range(world)]
        sent = {}
        for r in range(world):
            peers = [m[1] for m in msgs[r] if m[0] == 0]
            assert len(peers) == len(set(peers)) and r not in peers      # one message per peer
            for d_, peer, nbytes, tiles in msgs[r]:
                if d_ == 0:
                    sent[(r, peer)] = (nbytes, tiles)
        got = {}
        for d in range(world):
            for d_, peer, nbytes, tiles in msgs[d]:
                if d_ == 1:
                    got[(peer, d)] = (nbytes, tiles)
        assert sent == got
        want = {}
        for w, i, j, root, dests in plan_ops:
            if w == which:
                for d in dests:
                    want.setdefault((root, d), []).append((i, j))
        assert {key: v[1] for key, v in sent.items()} == want
        for (r, d), (nbytes, tiles) in sent.items():
            rows = lambda i: min(nb, n - i * nb)
            assert nbytes == 8 * sum(rows(i) * rows(j) for i, j in tiles)

    for k in (0, 3, nt - 2):
        ops = sl.plan_potrf_step(H, k)
        check("potrf", (H,), k, 0, ops)
        check("potrf", (H,), k, 1, ops)
        ops = sl.plan_gemm_step(A, B, Cm, k)
        check("gemm", (A, B, Cm), k, 0, ops)
        check("gemm", (A, B, Cm), k, 1, ops)
    sl.finalize()
