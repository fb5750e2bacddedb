"""Worker of tests/test_multirank_cpu.py: runs the N > 1 host logic of slate::gemm and
slate::potrf on CPU ranks.  The broadcast plans come from the product's own C++
(BaseMatrix::bcastPlan through the same list builders the drivers use); the bytes travel
over torch.distributed (gloo) instead of NCCL and the tile arithmetic is done by the CPU
oracle, so what is under test is ownership, destination sets and ordering: every tile a
rank needs must have been delivered to it, and the assembled result must equal the
oracle's single-process result.

usage: torchrun --nproc-per-node 2 tests/mp_worker.py P Q
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol  # noqa: E402
from slate_d35_b200 import host as sl  # noqa: E402

EPS = np.finfo(np.float64).eps


def tile_shape(m, n, nb, i, j):
    return min(nb, m - i * nb), min(nb, n - j * nb)


def run_bcasts(ops, which, tiles, rank, shape_of):
    """Execute the plan entries of one list in order: root sends, destinations receive."""
    for w, i, j, root, dests in ops:
        if w != which:
            continue
        if rank == root:
            t = torch.from_numpy(np.ascontiguousarray(tiles[(i, j)]))
            for d in dests:
                dist.send(t, d)
        elif rank in dests:
            buf = torch.empty(shape_of(i, j), dtype=torch.float64)
            dist.recv(buf, root)
            tiles[(i, j)] = np.asfortranarray(buf.numpy())


def gather_full(tiles, M, m, n, nb, rank, world):
    """rank 0 assembles the matrix from every rank's local tiles."""
    full = np.zeros((m, n), order="F")
    for i in range((m + nb - 1) // nb):
        for j in range((n + nb - 1) // nb):
            owner = sl.tile_rank(M, i, j)
            if M.kind != sl.GENERAL and i < j:
                continue
            shp = tile_shape(m, n, nb, i, j)
            if owner == 0:
                if rank == 0:
                    full[i * nb:i * nb + shp[0], j * nb:j * nb + shp[1]] = tiles[(i, j)]
            elif rank == owner:
                dist.send(torch.from_numpy(np.ascontiguousarray(tiles[(i, j)])), 0)
            elif rank == 0:
                buf = torch.empty(shp, dtype=torch.float64)
                dist.recv(buf, owner)
                full[i * nb:i * nb + shp[0], j * nb:j * nb + shp[1]] = buf.numpy()
    return full


def main():
    p, q = int(sys.argv[1]), int(sys.argv[2])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    assert p * q == world
    sl.initialize(0, rank, world, None)          # host-only: no GPU, no NCCL
    O = ol.oracle()

    # ------------------------------ gemm (SUMMA) ------------------------------
    m, n, k, nb = 44, 37, 29, 8                    # ragged in every dimension
    alpha, beta = 3.141592653589793, 2.718281828459045
    A = sl.Matrix(m, k, nb, p, q, origin="N")
    B = sl.Matrix(k, n, nb, p, q, origin="N")
    Cm = sl.Matrix(m, n, nb, p, q, origin="N")
    A0, B0, C0 = ol.matgen(1, m, k), ol.matgen(2, k, n), ol.matgen(3, m, n)
    mt, nt, kt = -(-m // nb), -(-n // nb), -(-k // nb)

    def local(M, full, rows, cols):
        out = {}
        for i in range(rows):
            for j in range(cols):
                if sl.tile_rank(M, i, j) == rank:
                    out[(i, j)] = np.asfortranarray(full[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb].copy())
        return out

    a_t, b_t, c_t = local(A, A0, mt, kt), local(B, B0, kt, nt), local(Cm, C0, mt, nt)
    for kk in range(kt):
        ops = sl.plan_gemm_step(A, B, Cm, kk)
        run_bcasts(ops, 0, a_t, rank, lambda i, j: tile_shape(m, k, nb, i, j))
        run_bcasts(ops, 1, b_t, rank, lambda i, j: tile_shape(k, n, nb, i, j))
        for (i, j), c in c_t.items():
            a, b = a_t[(i, kk)], b_t[(kk, j)]          # KeyError = a needed tile was not delivered
            O.oracle_tile_dgemm(ord("N"), ord("N"), c.shape[0], c.shape[1], a.shape[1], alpha,
                                ol.P(a), a.shape[0], ol.P(b), b.shape[0], beta if kk == 0 else 1.0,
                                ol.P(c), c.shape[0])
        # release remote panel tiles, as the driver does
        for key in [t for t in a_t if t[1] == kk and sl.tile_rank(A, *t) != rank]:
            del a_t[key]
        for key in [t for t in b_t if t[0] == kk and sl.tile_rank(B, *t) != rank]:
            del b_t[key]
    got = gather_full(c_t, Cm, m, n, nb, rank, world)
    if rank == 0:
        want = C0.copy(order="F")
        O.oracle_dgemm(ord("N"), ord("N"), m, n, k, nb, alpha, ol.P(A0), m, ol.P(B0), k, beta, ol.P(want), m)
        assert np.array_equal(got, want), np.abs(got - want).max()   # same tile arithmetic: bit exact

    # ------------------------------ gemmA (A stationary, C reduced) ------------------------------
    # src/gemmA.cc: B(:, kc) travels to the owners of the block columns of A, every rank
    # multiplies its local tiles of A, partial tiles of C(:, kc) are summed onto their owners.
    n2 = 13                                        # two block columns of B and C, the last ragged
    B2 = sl.Matrix(k, n2, nb, p, q, origin="N")
    C2 = sl.Matrix(m, n2, nb, p, q, origin="N")
    B20, C20 = ol.matgen(4, k, n2), ol.matgen(5, m, n2)
    nt2 = -(-n2 // nb)
    b2_t, c2_t = local(B2, B20, kt, nt2), local(C2, C20, mt, nt2)
    for kc in range(nt2):
        ops = sl.plan_gemmA_step(A, B2, C2, kc)
        run_bcasts(ops, 0, b2_t, rank, lambda i, j: tile_shape(k, n2, nb, i, j))
        partial = {}
        for i in range(mt):
            mine = [j for j in range(kt) if sl.tile_rank(A, i, j) == rank]
            owner = sl.tile_rank(C2, i, kc) == rank
            if owner:
                partial[i] = beta * c2_t[(i, kc)]
            elif mine:
                partial[i] = np.zeros(tile_shape(m, n2, nb, i, kc), order="F")
            for j in mine:
                partial[i] += alpha * a_t[(i, j)] @ b2_t[(j, kc)]   # KeyError = B tile not delivered
        for w, i, j, root, sources in ops:
            if w != 2:
                continue
            assert j == kc
            if rank == root:
                for src in sources:
                    buf = torch.empty(tile_shape(m, n2, nb, i, kc), dtype=torch.float64)
                    dist.recv(buf, src)
                    partial[i] = partial[i] + buf.numpy()
            elif rank in sources:
                dist.send(torch.from_numpy(np.ascontiguousarray(partial.pop(i))), root)
        for i, t in partial.items():
            # whatever is left must be an owned tile: a partial sum nobody collected is a bug
            assert sl.tile_rank(C2, i, kc) == rank, (i, kc)
            c2_t[(i, kc)] = np.asfortranarray(t)
        for key in [t for t in b2_t if t[1] == kc and sl.tile_rank(B2, *t) != rank]:
            del b2_t[key]
    got = gather_full(c2_t, C2, m, n2, nb, rank, world)
    if rank == 0:
        want = alpha * A0 @ B20 + beta * C20
        assert np.abs(got - want).max() <= 3 * np.sqrt(k) * EPS * np.abs(want).max()

    # ------------------------------ getrf_nopiv ------------------------------
    # src/getrf_nopiv.cc: diagonal tile factored by its owner and sent down its column and along
    # its row; column tiles solved against U and sent along their rows; row tiles solved
    # against L and sent down their columns; every rank updates its own trailing tiles.
    def lu_nopiv(a):
        a = a.copy()
        for c in range(min(a.shape)):
            a[c + 1:, c] /= a[c, c]
            a[c + 1:, c + 1:] -= np.outer(a[c + 1:, c], a[c, c + 1:])
        return a

    ml, nl, nbl = 45, 38, 8
    G = sl.Matrix(ml, nl, nbl, p, q, origin="N")
    G0 = ol.matgen(9, ml, nl)
    G0[np.arange(min(ml, nl)), np.arange(min(ml, nl))] += max(ml, nl)
    mtl, ntl = -(-ml // nbl), -(-nl // nbl)
    g_t = {}
    for i in range(mtl):
        for j in range(ntl):
            if sl.tile_rank(G, i, j) == rank:
                g_t[(i, j)] = np.asfortranarray(G0[i * nbl:(i + 1) * nbl, j * nbl:(j + 1) * nbl].copy())
    shape_g = lambda i, j: tile_shape(ml, nl, nbl, i, j)
    for kk in range(min(mtl, ntl)):
        ops = sl.plan_getrf_nopiv_step(G, kk)
        if sl.tile_rank(G, kk, kk) == rank:
            g_t[(kk, kk)] = np.asfortranarray(lu_nopiv(g_t[(kk, kk)]))
        run_bcasts(ops, 0, g_t, rank, shape_g)
        for i in range(kk + 1, mtl):                        # A(i,k) U^-1
            if sl.tile_rank(G, i, kk) == rank:
                d = g_t[(kk, kk)]
                U = np.triu(d[:, :d.shape[1]])[:d.shape[1], :]
                g_t[(i, kk)] = np.asfortranarray(np.linalg.solve(U.T, g_t[(i, kk)].T).T)
        run_bcasts(ops, 1, g_t, rank, shape_g)
        for j in range(kk + 1, ntl):                        # L^-1 A(k,j)
            if sl.tile_rank(G, kk, j) == rank:
                d = g_t[(kk, kk)]
                Lk = np.tril(d[:, :d.shape[0]], -1)[:, :d.shape[0]] + np.eye(d.shape[0])
                g_t[(kk, j)] = np.asfortranarray(np.linalg.solve(Lk, g_t[(kk, j)]))
        run_bcasts(ops, 2, g_t, rank, shape_g)
        for i in range(kk + 1, mtl):
            for j in range(kk + 1, ntl):
                if sl.tile_rank(G, i, j) == rank:
                    g_t[(i, j)] = g_t[(i, j)] - g_t[(i, kk)] @ g_t[(kk, j)]   # KeyError = not delivered
        for key in [t for t in g_t if (t[0] == kk or t[1] == kk) and sl.tile_rank(G, *t) != rank]:
            del g_t[key]
    got = gather_full(g_t, G, ml, nl, nbl, rank, world)
    if rank == 0:
        mn = min(ml, nl)
        Lf = np.tril(got[:, :mn], -1) + np.eye(ml, mn)
        Uf = np.triu(got[:mn, :])
        assert np.abs(Lf @ Uf - G0).max() <= 50 * EPS * max(ml, nl) * np.abs(G0).max()

    # ------------------------------ hemm (side Left, lower storage) ------------------------------
    # src/hemmC.cc: block column k of the Hermitian A is read from the stored triangle where it
    # exists and as the transposed block row where it does not; both travel to the owners of
    # the block rows of C.
    nh, nn, nbh = 43, 21, 8
    Hh = sl.HermitianMatrix("L", nh, nbh, p, q, origin="N")
    Bh = sl.Matrix(nh, nn, nbh, p, q, origin="N")
    Ch = sl.Matrix(nh, nn, nbh, p, q, origin="N")
    S0 = ol.matgen(10, nh, nh)
    S0 = (S0 + S0.T) / 2
    Bh0, Ch0 = ol.matgen(11, nh, nn), ol.matgen(12, nh, nn)
    mth, nth = -(-nh // nbh), -(-nn // nbh)
    s_t = {(i, j): np.asfortranarray(S0[i * nbh:(i + 1) * nbh, j * nbh:(j + 1) * nbh].copy())
           for i in range(mth) for j in range(i + 1) if sl.tile_rank(Hh, i, j) == rank}
    bh_t = {(i, j): np.asfortranarray(Bh0[i * nbh:(i + 1) * nbh, j * nbh:(j + 1) * nbh].copy())
            for i in range(mth) for j in range(nth) if sl.tile_rank(Bh, i, j) == rank}
    ch_t = {(i, j): np.asfortranarray(Ch0[i * nbh:(i + 1) * nbh, j * nbh:(j + 1) * nbh].copy())
            for i in range(mth) for j in range(nth) if sl.tile_rank(Ch, i, j) == rank}
    for kk in range(mth):
        ops = sl.plan_hemm_step(Hh, Bh, Ch, kk)
        run_bcasts(ops, 0, s_t, rank, lambda i, j: tile_shape(nh, nh, nbh, i, j))
        run_bcasts(ops, 1, bh_t, rank, lambda i, j: tile_shape(nh, nn, nbh, i, j))
        for (i, j), c in ch_t.items():
            if i == kk:
                d = s_t[(kk, kk)]
                a_ik = np.tril(d) + np.tril(d, -1).T
            elif i > kk:
                a_ik = s_t[(i, kk)]
            else:
                a_ik = s_t[(kk, i)].T
            ch_t[(i, j)] = alpha * a_ik @ bh_t[(kk, j)] + (beta if kk == 0 else 1.0) * c
        for key in [t for t in s_t if (t[1] == kk or t[0] == kk) and sl.tile_rank(Hh, *t) != rank]:
            del s_t[key]
        for key in [t for t in bh_t if t[0] == kk and sl.tile_rank(Bh, *t) != rank]:
            del bh_t[key]
    got = gather_full(ch_t, Ch, nh, nn, nbh, rank, world)
    if rank == 0:
        want = alpha * S0 @ Bh0 + beta * Ch0
        assert np.abs(got - want).max() <= 3 * np.sqrt(nh) * EPS * np.abs(want).max()

    # ------------------------------ potrf ------------------------------
    n, nb = 52, 8
    H = sl.HermitianMatrix("L", n, nb, p, q, origin="N")
    H0 = ol.matgen(7, n, n)
    H0[np.arange(n), np.arange(n)] += n
    nt = -(-n // nb)
    h_t = {}
    for i in range(nt):
        for j in range(i + 1):
            if sl.tile_rank(H, i, j) == rank:
                h_t[(i, j)] = np.asfortranarray(H0[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb].copy())
    info = 0
    for kk in range(nt):
        ops = sl.plan_potrf_step(H, kk)
        if sl.tile_rank(H, kk, kk) == rank:
            t = h_t[(kk, kk)]
            iinfo = O.oracle_tile_dpotrf(ord("L"), t.shape[0], ol.P(t), t.shape[0])
            if iinfo and not info:
                info = kk * nb + iinfo
        run_bcasts(ops, 0, h_t, rank, lambda i, j: tile_shape(n, n, nb, i, j))
        for i in range(kk + 1, nt):
            if sl.tile_rank(H, i, kk) == rank:
                t, l = h_t[(i, kk)], h_t[(kk, kk)]
                O.oracle_tile_dtrsm(ord("R"), ord("L"), ord("C"), ord("N"), t.shape[0], t.shape[1], 1.0,
                                    ol.P(l), l.shape[0], ol.P(t), t.shape[0])
        run_bcasts(ops, 1, h_t, rank, lambda i, j: tile_shape(n, n, nb, i, j))
        for j in range(kk + 1, nt):
            for i in range(j, nt):
                if sl.tile_rank(H, i, j) != rank:
                    continue
                c, ai, aj = h_t[(i, j)], h_t[(i, kk)], h_t[(j, kk)]
                if i == j:
                    O.oracle_tile_dsyrk(ord("L"), ord("N"), c.shape[0], ai.shape[1], -1.0, ol.P(ai),
                                        ai.shape[0], 1.0, ol.P(c), c.shape[0])
                else:
                    O.oracle_tile_dgemm(ord("N"), ord("C"), c.shape[0], c.shape[1], ai.shape[1], -1.0,
                                        ol.P(ai), ai.shape[0], ol.P(aj), aj.shape[0], 1.0, ol.P(c),
                                        c.shape[0])
        for key in [t for t in h_t if t[1] == kk and sl.tile_rank(H, *t) != rank]:
            del h_t[key]
    got = gather_full(h_t, H, n, n, nb, rank, world)
    if rank == 0:
        want = H0.copy(order="F")
        assert O.oracle_dpotrf(ord("L"), n, nb, ol.P(want), n) == 0 and info == 0
        assert np.array_equal(np.tril(got), np.tril(want)), np.abs(np.tril(got) - np.tril(want)).max()
        print(f"MP_CPU_OK {p}x{q}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
