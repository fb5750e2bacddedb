"""Multi-rank parity: slate::gemm and slate::potrf on a p x q grid (one rank per GPU, NCCL
tile broadcasts) against the CPU oracle.  Run under torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29511 scripts/mp_parity.py

Every rank gathers its own tiles (zeros elsewhere); an all-reduce sum assembles the matrix.
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

GRIDS = {1: [(1, 1)], 2: [(1, 2), (2, 1)], 4: [(2, 2), (1, 4)], 8: [(2, 4), (4, 2)]}
EPS = np.finfo(np.float64).eps


def assemble(M):
    t = torch.from_numpy(np.ascontiguousarray(M.to_numpy())).cuda()
    dist.all_reduce(t)
    return np.asfortranarray(t.cpu().numpy())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    lr = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    obj = [sl.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    sl.initialize(lr, rank, world, obj[0])
    fails = 0
    for (p, q) in GRIDS[world]:
        for la in (0, 1, 2):
            # ---- gemm, ragged tiles, tester alpha/beta ----
            m, n, k, nb = 1500, 1300, 1100, 256
            alpha, beta = 3.141592653589793, 2.718281828459045
            A = sl.Matrix(m, k, nb, p, q).generate(1)
            B = sl.Matrix(k, n, nb, p, q).generate(2)
            Cm = sl.Matrix(m, n, nb, p, q).generate(3)
            sl.gemm(alpha, A, B, beta, Cm, lookahead=la)
            got = assemble(Cm)
            if rank == 0:
                A0, B0, want = ol.matgen(1, m, k), ol.matgen(2, k, n), ol.matgen(3, m, n)
                ol.oracle().oracle_dgemm(ord("N"), ord("N"), m, n, k, nb, alpha, ol.P(A0), m, ol.P(B0),
                                         k, beta, ol.P(want), m)
                err = np.abs(got - want).max() / np.abs(want).max()
                ok = err <= 3 * np.sqrt(k) * EPS
                fails += (not ok)
                print(f"gemm  {p}x{q} la={la}: err {err:.2e} {'ok' if ok else 'FAIL'}", flush=True)
            for M in (A, B, Cm):
                M.free()
            # ---- potrf ----
            for uplo in "LU":
                n, nb = 1800, 256
                H = sl.HermitianMatrix(uplo, n, nb, p, q).generate(7, dominant=True)
                info = sl.potrf(H, lookahead=la)
                got = assemble(H)
                berr_dev = None
                if uplo == "L":
                    # the engine's own distributed residual (what bench.py uses at full size)
                    r, a0 = sl.potrf_residual(H, 7, True)
                    t = torch.tensor([r, a0], dtype=torch.float64, device="cuda")
                    dist.all_reduce(t)
                    berr_dev = float(t[0].sqrt() / (n * t[1].sqrt()))
                if rank == 0:
                    want = ol.matgen(7, n, n)
                    want[np.arange(n), np.arange(n)] += n
                    tri = np.tril if uplo == "L" else np.triu
                    # dense backward error of the assembled factor, both triangles
                    S = tri(want) + tri(want, -1 if uplo == "L" else 1).T
                    F = tri(got)
                    R = (F @ F.T if uplo == "L" else F.T @ F) - S
                    berr = float(np.linalg.norm(R) / (n * np.linalg.norm(S)))
                    if berr_dev is not None:
                        berr = max(berr, berr_dev)
                    assert ol.oracle().oracle_dpotrf(ord(uplo), n, nb, ol.P(want), n) == 0
                    err = np.abs(tri(got) - tri(want)).max() / np.abs(tri(want)).max()
                    ok = info == 0 and err <= 50 * EPS and berr <= 50 * EPS
                    fails += (not ok)
                    print(f"potrf {p}x{q} la={la} {uplo}: info {info} err {err:.2e} backward {berr:.2e} "
                          f"{'ok' if ok else 'FAIL'}", flush=True)
                H.free()
    # herk and posv drivers on every grid: the tester's solve residual
    for (p, q) in GRIDS[world]:
        n, nb, nrhs, kk = 1400, 256, 10, 700
        H = sl.HermitianMatrix("L", n, nb, p, q).generate(7, dominant=True)
        B = sl.Matrix(n, nrhs, nb, p, q).generate(8)
        info = sl.posv(H, B)
        X = assemble(B)
        Ak = sl.Matrix(n, kk, nb, p, q).generate(9, dist=2)
        Ck = sl.HermitianMatrix("L", n, nb, p, q).generate(10, dist=2)
        sl.herk(-1.0, Ak, 1.0, Ck)
        Cgot = assemble(Ck)
        if rank == 0:
            A0 = ol.matgen(7, n, n)
            A0[np.arange(n), np.arange(n)] += n
            S = np.tril(A0) + np.tril(A0, -1).T
            B0 = ol.matgen(8, n, nrhs)
            resid = np.linalg.norm(B0 - S @ X, 1) / (n * np.linalg.norm(S, 1) * np.linalg.norm(X, 1))
            Ak0, Ck0 = ol.matgen(9, n, kk, dist=2), ol.matgen(10, n, n, dist=2)
            want = np.tril(-Ak0 @ Ak0.T + Ck0)
            herr = np.abs(np.tril(Cgot) - want).max() / np.abs(want).max()
            ok = info == 0 and resid <= 25 * EPS and herr <= 3 * np.sqrt(kk) * EPS
            fails += (not ok)
            print(f"posv  {p}x{q}: info {info} residual {resid:.2e}; herk err {herr:.2e} {'ok' if ok else 'FAIL'}", flush=True)
        for M in (H, B, Ak, Ck):
            M.free()

    # host-origin matrices in ScaLAPACK layout (the e2e path of bench.py): every rank passes its
    # own 2D block-cyclic local array, tiles are staged H2D / D2H around the broadcasts
    for (p, q) in GRIDS[world]:
        m, n, k, nb = 1500, 1300, 1100, 256
        myrow, mycol = rank % p, rank // p

        def local_of(full, rows, cols):
            ri = [i for t in range(myrow, -(-rows // nb), p) for i in range(t * nb, min((t + 1) * nb, rows))]
            ci = [j for t in range(mycol, -(-cols // nb), q) for j in range(t * nb, min((t + 1) * nb, cols))]
            return np.asfortranarray(full[np.ix_(ri, ci)]), ri, ci

        A0, B0, C0 = ol.matgen(1, m, k), ol.matgen(2, k, n), ol.matgen(3, m, n)
        Al, _, _ = local_of(A0, m, k)
        Bl, _, _ = local_of(B0, k, n)
        Cl, ri, ci = local_of(C0, m, n)
        if Al.size and Bl.size and Cl.size:
            A = sl.Matrix.fromScaLAPACK(m, k, Al, nb, p, q)
            B = sl.Matrix.fromScaLAPACK(k, n, Bl, nb, p, q)
            Cm = sl.Matrix.fromScaLAPACK(m, n, Cl, nb, p, q)
            sl.gemm(3.141592653589793, A, B, 2.718281828459045, Cm)
            full = np.zeros((m, n))
            full[np.ix_(ri, ci)] = Cl
            t = torch.from_numpy(full).cuda()
            dist.all_reduce(t)
            if rank == 0:
                want = C0.copy(order="F")
                ol.oracle().oracle_dgemm(ord("N"), ord("N"), m, n, k, nb, 3.141592653589793, ol.P(A0), m,
                                         ol.P(B0), k, 2.718281828459045, ol.P(want), m)
                err = np.abs(t.cpu().numpy() - want).max() / np.abs(want).max()
                ok = err <= 3 * np.sqrt(k) * EPS
                fails += (not ok)
                print(f"gemm  {p}x{q} fromScaLAPACK host buffers: err {err:.2e} {'ok' if ok else 'FAIL'}", flush=True)
            for M in (A, B, Cm):
                M.free()

    # gemmA (A stationary, B broadcast, partial C tiles reduced over NCCL): two block columns
    # of B, transposed A, beta != 1; and the one-block-column case slate::gemm picks by itself
    for (p, q) in GRIDS[world]:
        m, k, nb = 1500, 1100, 256
        for (n, opa, method) in ((400, "N", "A"), (400, "T", "A"), (10, "N", "auto")):
            alpha, beta = 3.141592653589793, 2.718281828459045
            ash = (m, k) if opa == "N" else (k, m)
            A = sl.Matrix(*ash, nb, p, q).generate(1)
            B = sl.Matrix(k, n, nb, p, q).generate(2)
            Cm = sl.Matrix(m, n, nb, p, q).generate(3)
            sl.gemm(alpha, A, B, beta, Cm, opa=opa, method=method)
            got = assemble(Cm)
            if rank == 0:
                A0, B0, C0 = ol.matgen(1, *ash), ol.matgen(2, k, n), ol.matgen(3, m, n)
                want = alpha * (A0 if opa == "N" else A0.T) @ B0 + beta * C0
                err = np.abs(got - want).max() / np.abs(want).max()
                ok = err <= 3 * np.sqrt(k) * EPS
                fails += (not ok)
                print(f"gemmA {p}x{q} n={n} op(A)={opa} method={method}: err {err:.2e} {'ok' if ok else 'FAIL'}", flush=True)
            for M in (A, B, Cm):
                M.free()

    # hemm / symm, her2k and getrf_nopiv on the grids (their broadcast plans run over gloo in
    # tests/mp_worker.py and are replayed on 2x2 / 2x4 in tests/test_plans_grid.py; this is the
    # same on GPUs).  MP_PARITY_EXTRA=0 skips it.
    for (p, q) in (GRIDS[world] if os.environ.get("MP_PARITY_EXTRA", "1") == "1" else []):
        n, nrhs, nb = 1500, 700, 256
        S0 = ol.matgen(30, n, n)
        S0 = (S0 + S0.T) / 2
        B0, C0 = ol.matgen(31, n, nrhs), ol.matgen(32, n, nrhs)
        H = sl.HermitianMatrix("L", n, nb, p, q).from_numpy(np.tril(S0))
        B = sl.Matrix(n, nrhs, nb, p, q).from_numpy(B0)
        Cm = sl.Matrix(n, nrhs, nb, p, q).from_numpy(C0)
        sl.hemm("L", 3.0, H, B, -0.5, Cm)
        got = assemble(Cm)
        herr = np.abs(got - (3.0 * S0 @ B0 - 0.5 * C0)).max() / np.abs(S0 @ B0).max()

        K0, L0 = ol.matgen(33, n, nrhs), ol.matgen(34, n, nrhs)
        Ak = sl.Matrix(n, nrhs, nb, p, q).from_numpy(K0)
        Bk = sl.Matrix(n, nrhs, nb, p, q).from_numpy(L0)
        Ck = sl.HermitianMatrix("L", n, nb, p, q).from_numpy(np.tril(S0))
        sl.her2k(2.0, Ak, Bk, 1.0, Ck)
        got = assemble(Ck)
        want = 2.0 * (K0 @ L0.T + L0 @ K0.T) + S0
        kerr = np.abs(np.tril(got) - np.tril(want)).max() / np.abs(want).max()

        G0 = ol.matgen(35, n, n)
        G0[np.arange(n), np.arange(n)] += n
        G = sl.Matrix(n, n, nb, p, q).from_numpy(G0)
        info = sl.getrf_nopiv(G)
        got = assemble(G)
        Lf, Uf = np.tril(got, -1) + np.eye(n), np.triu(got)
        lerr = np.linalg.norm(Lf @ Uf - G0) / (n * np.linalg.norm(G0))
        if rank == 0:
            ok = herr <= 3 * np.sqrt(n) * EPS and kerr <= 3 * np.sqrt(2 * nrhs) * EPS and info == 0 and lerr <= 50 * EPS
            fails += (not ok)
            print(f"hemm/her2k/getrf_nopiv {p}x{q}: hemm err {herr:.2e}, her2k err {kerr:.2e}, "
                  f"LU info {info} backward {lerr:.2e} {'ok' if ok else 'FAIL'}", flush=True)
        for M in (H, B, Cm, Ak, Bk, Ck, G):
            M.free()

    # the ScaLAPACK-convention entry points (include/slate_b200_scalapack.h) as an unmodified
    # ScaLAPACK program calls them: pdgemm (with transposes) and pdposv on both rank orders
    from slate_d35_b200 import scalapack as sc
    for gi, (p, q) in enumerate(GRIDS[world]):
        for order in "CR":
            ctxt = 10 + 2 * gi + (order == "R")
            sc.grid_set(ctxt, order, p, q)
            myrow, mycol = (rank % p, rank // p) if order == "C" else (rank // q, rank % q)
            m, n, k, nb, nrhs = 1500, 1300, 1100, 256, 10

            def gather(local, ri, ci, shape):
                full = np.zeros(shape)
                full[np.ix_(ri, ci)] = local
                t = torch.from_numpy(full).cuda()
                dist.all_reduce(t)
                return t.cpu().numpy()

            A0, B0, C0 = ol.matgen(1, k, m), ol.matgen(2, k, n), ol.matgen(3, m, n)
            Al, _, _ = sc.local_part(A0, nb, nb, myrow, mycol, p, q)
            Bl, _, _ = sc.local_part(B0, nb, nb, myrow, mycol, p, q)
            Cl, ri, ci = sc.local_part(C0, nb, nb, myrow, mycol, p, q)
            sc.pgemm("T", "N", m, n, k, 3.141592653589793, Al, 1, 1, sc.descinit(k, m, nb, nb, ctxt, Al.shape[0]),
                     Bl, 1, 1, sc.descinit(k, n, nb, nb, ctxt, Bl.shape[0]), 2.718281828459045,
                     Cl, 1, 1, sc.descinit(m, n, nb, nb, ctxt, Cl.shape[0]))
            got = gather(Cl, ri, ci, (m, n))
            want = 3.141592653589793 * A0.T @ B0 + 2.718281828459045 * C0
            gerr = np.abs(got - want).max() / np.abs(want).max()

            nn = 1536
            S = ol.matgen(7, nn, nn)
            S = (S + S.T) / 2
            S[np.arange(nn), np.arange(nn)] += nn
            R0 = ol.matgen(8, nn, nrhs)
            Sl, _, _ = sc.local_part(np.tril(S), nb, nb, myrow, mycol, p, q)
            Rl, ri, ci = sc.local_part(R0, nb, nb, myrow, mycol, p, q)
            info = sc.pposv("L", nn, nrhs, Sl, 1, 1, sc.descinit(nn, nn, nb, nb, ctxt, Sl.shape[0]),
                            Rl, 1, 1, sc.descinit(nn, nrhs, nb, nb, ctxt, max(1, Rl.shape[0])))
            X = gather(Rl, ri, ci, (nn, nrhs))
            resid = np.linalg.norm(R0 - S @ X, 1) / (nn * np.linalg.norm(S, 1) * np.linalg.norm(X, 1))
            if rank == 0:
                ok = gerr <= 3 * np.sqrt(k) * EPS and info == 0 and resid <= 25 * EPS
                fails += (not ok)
                print(f"scalapack api {p}x{q} order {order}: pdgemm(T,N) err {gerr:.2e}, pdposv info {info} "
                      f"residual {resid:.2e} {'ok' if ok else 'FAIL'}", flush=True)

    # info must be the same on every rank
    n, nb = 1200, 256
    p, q = GRIDS[world][0]
    A0 = ol.matgen(7, n, n)
    A0[np.arange(n), np.arange(n)] += n
    A0[700, 700] = -3.0
    H = sl.HermitianMatrix("L", n, nb, p, q).from_numpy(A0)
    info = sl.potrf(H)
    t = torch.tensor([info], dtype=torch.int64, device="cuda")
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if rank == 0:
        ok = int(lo.item()) == int(hi.item()) == 701
        fails += (not ok)
        print(f"potrf info on all ranks: {int(lo.item())}..{int(hi.item())} {'ok' if ok else 'FAIL'}", flush=True)
    H.free()
    sl.finalize()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MP_PARITY", "PASS" if fails == 0 else f"FAIL ({fails})", flush=True)
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
