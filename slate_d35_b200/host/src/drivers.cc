// drivers.cc -- slate::gemm (gemmC SUMMA) and slate::potrf on the tile engine, plus the
// harness helpers declared in slate/slate.hh.
//
// The loop structure follows the reference drivers (src/gemmC.cc:29-191,
// src/potrf.cc:23-210); what differs is the execution model.  The reference builds an
// OpenMP task DAG whose tasks block in queue->sync(); here ONE host thread enqueues every
// step on CUDA streams and never waits: dependencies between the panel queue, the
// lookahead queues, the trailing-update queue and the NCCL broadcasts are events derived
// from which queue last wrote each tile (BaseMatrix::tileGet), and broadcast workspace is
// recycled by deferred release.  Lookahead keeps its meaning: how many block columns the
// broadcasts / panel factorisations run ahead of the trailing update.
#include <deque>
#include <thread>
#include "slate/slate.hh"

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>

namespace slate {

namespace {
/// SLATE_B200_TIMING=1: print where the enqueueing host thread spends its time.
struct HostTimer {
    bool on = std::getenv("SLATE_B200_TIMING") != nullptr;
    double acc[8] = {};
    std::chrono::steady_clock::time_point t0;
    void start() { if (on) t0 = std::chrono::steady_clock::now(); }
    void stop(int slot)
    {
        if (on) acc[slot] += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
};
}  // namespace

//------------------------------------------------------------------------------
template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag gemm_bcast_list(int which, Matrix<scalar_t>& A,
                                                        Matrix<scalar_t>& B, Matrix<scalar_t>& C,
                                                        int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    if (which == 0) {
        for (int64_t i = 0; i < A.mt(); ++i)
            list.push_back({i, k, {C.sub(i, i, 0, C.nt() - 1)}, i});
    }
    else {
        for (int64_t j = 0; j < B.nt(); ++j)
            list.push_back({k, j, {C.sub(0, C.mt() - 1, j, j)}, j});
    }
    return list;
}

template <typename scalar_t>
typename Matrix<scalar_t>::BcastListTag potrf_bcast_list(int which, HermitianMatrix<scalar_t>& A,
                                                         int64_t k)
{
    typename Matrix<scalar_t>::BcastListTag list;
    const int64_t A_nt = A.nt();
    if (which == 0) {
        if (k + 1 <= A_nt - 1)
            list.push_back({k, k, {A.sub(k + 1, A_nt - 1, k, k)}, k});
    }
    else {
        for (int64_t i = k + 1; i < A_nt; ++i)
            list.push_back({i, k, {A.sub(i, i, k + 1, i), A.sub(i, A_nt - 1, i, i)}, i});
    }
    return list;
}

//------------------------------------------------------------------------------
template <typename scalar_t>
void gemmC(scalar_t alpha, Matrix<scalar_t>& A, Matrix<scalar_t>& B, scalar_t beta,
           Matrix<scalar_t>& C, Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::gemm: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const Layout layout = Layout::ColMajor;
    const scalar_t one = 1.0;
    const int queue_0 = 0;
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;

    slate_assert(A.mt() == C.mt());
    slate_assert(B.nt() == C.nt());
    slate_assert(A.nt() == B.mt());
    slate_assert(C.op() == Op::NoTrans);
    const int64_t kt = A.nt();

    // src/gemmC.cc:57-60
    C.allocateBatchArrays();
    C.reserveDeviceWorkspace(0);

    // Device-origin C: one launch per block step, as the reference.
    // Host-origin C (tiles staged over PCIe): C is cut into NSLAB slabs of block columns and
    // slab s runs its k loop `skew` steps behind slab s-1, all slabs that are active in an
    // iteration sharing one launch.  Slab 0 starts as soon as it has arrived and the others
    // are fetched behind the compute of the earlier ones; at the end slab s is written back
    // while the later slabs still compute.  The reference fetches all of C before the first
    // step and copies it back after the last (src/gemmC.cc:70-74, 187-190), which leaves
    // both transfers exposed (2 x 0.7 s for n = 65536 on one GPU).
    const bool staged = (C.origin() != Target::Devices) && C.num_devices() > 0;
    const int64_t nslab = staged ? std::max<int64_t>(1, std::min<int64_t>(8, C.nt())) : 1;
    // origin tiles travelling over PCIe during the k loop: descriptors must not queue on the
    // copy engine behind them (Engine::upload_desc)
    struct BulkTrafficGuard {
        explicit BulkTrafficGuard(bool on) { Engine::instance().set_bulk_host_traffic(on); }
        ~BulkTrafficGuard() { Engine::instance().set_bulk_host_traffic(false); }
    } bulk_guard(A.origin() != Target::Devices || B.origin() != Target::Devices
                 || C.origin() != Target::Devices);
    // the first and the last slab are half as wide as the others: the first one's transfer
    // is the start-up time nothing can hide, the last one's write-back is the tail
    std::vector<int64_t> slab_begin(nslab + 1, 0);
    {
        const bool weighted = nslab > 2 && C.nt() >= 4 * nslab;         // weights 1, 2, ..., 2, 1
        const int64_t units = weighted ? 2 * nslab - 2 : nslab;
        int64_t acc = 0;
        for (int64_t s = 0; s < nslab; ++s) {
            slab_begin[s] = acc * C.nt() / units;
            acc += (weighted && s > 0 && s < nslab - 1) ? 2 : 1;
        }
        slab_begin[nslab] = C.nt();
    }
    auto slab_cols = [&](int64_t s, int64_t* j1, int64_t* j2) {
        *j1 = slab_begin[s];
        *j2 = slab_begin[s + 1] - 1;
    };
    int64_t skew = 0;
    if (nslab > 1) {
        // steps of compute that cover the transfer of all of C: local bytes over ~45 GB/s of
        // PCIe against local flops per step at ~30 TFLOP/s
        double local_tiles = (double)C.getMaxDeviceTiles(0);
        double bytes = local_tiles * sizeof(scalar_t) * C.tileMb(0) * C.tileNb(0);
        double flops_step = local_tiles * 2.0 * C.tileMb(0) * C.tileNb(0) * A.tileNb(0)
                            * (is_complex<scalar_t>::value ? 4.0 : 1.0);
        double steps = (bytes / 45e9) / (flops_step / 30e12);
        skew = std::max<int64_t>(1, std::min<int64_t>((int64_t)std::ceil(steps * 1.25), 16));
    }
    const int64_t extra_panels = skew * (nslab - 1);

    // src/gemmC.cc:57-60 (panels stay resident until the last slab has used them)
    A.ensureDeviceWorkspace((lookahead + 3 + extra_panels) * A.mt());
    B.ensureDeviceWorkspace((lookahead + 3 + extra_panels) * B.nt());

    if (kt == 0) {
        // nothing to accumulate; the reference's loop does not run either
        C.tileGetAllForWritingOnDevices(LayoutConvert(layout), queue_0);
        C.tileUpdateAllOrigin();
        Engine::instance().sync_all();
        return;
    }

    // send block column k of A along the process rows and block row k of B along the
    // process columns that own C (src/gemmC.cc:78-110, 143-157)
    auto bcast = [&](int64_t k) {
        BcastListTag list_A = gemm_bcast_list(0, A, B, C, k);
        A.template listBcastMT<Target::Devices>(list_A, layout);
        BcastListTag list_B = gemm_bcast_list(1, A, B, C, k);
        B.template listBcastMT<Target::Devices>(list_B, layout);
    };
    // fetch_C (src/gemmC.cc:70-74), by block columns; the consuming launch waits for its own
    // tiles.  Slab s is first used in iteration skew * s; its columns are asked for a few per
    // iteration during the slab period before that, so that neither the copy engine (panels
    // queue behind them) nor the host (2048 copies take ~40 ms to enqueue) sees a burst.
    int64_t next_col = 0;
    auto fetch_cols_upto = [&](int64_t j_end) {        // columns [next_col, j_end)
        j_end = std::min<int64_t>(j_end, C.nt());
        if (j_end <= next_col) return;
        C.sub(0, C.mt() - 1, next_col, j_end - 1).tileGetAllForWritingOnDevices(LayoutConvert(layout), -1);
        next_col = j_end;
    };
    auto fetch_for_iteration = [&](int64_t t) {
        if (nslab == 1) return;
        const int64_t s = t / skew + 1;                  // the slab being prefetched
        if (s >= nslab) { fetch_cols_upto(C.nt()); return; }
        const int64_t r = t % skew;                       // complete one iteration early
        const int64_t parts = std::max<int64_t>(1, skew - 1);
        const int64_t width = slab_begin[s + 1] - slab_begin[s];
        fetch_cols_upto(slab_begin[s] + std::min<int64_t>(width, (width * (r + 1) + parts - 1) / parts));
    };

    // Write-back of finished slabs.  Its device-to-host copies wait (on the home stream) for the
    // slab's last launch, which the host has enqueued several launches ahead of the device; a
    // stream that is blocked on an event accepts only ~1000 more operations before the ENQUEUE
    // blocks the host (measured: the host then sat until the device had caught up, the device
    // queue ran dry and every slab cost 57 ms of idle GPU).  So a finished slab is queued here
    // and its copies are enqueued, a few block columns per iteration, once that launch is done.
    struct Writeback { int64_t j_next, j_end; uint64_t seq; };
    std::deque<Writeback> writebacks;
    auto drain_writebacks = [&](int64_t max_cols) {
        Engine& eng = Engine::instance();
        while (!writebacks.empty() && max_cols > 0) {
            Writeback& w = writebacks.front();
            if (!eng.queue_mark_done(queue_0, w.seq)) break;
            const int64_t j2 = std::min<int64_t>(w.j_end, w.j_next + max_cols);
            C.sub(0, C.mt() - 1, w.j_next, j2 - 1).tileUpdateAllOriginAsync();
            max_cols -= j2 - w.j_next;
            w.j_next = j2;
            if (w.j_next >= w.j_end) writebacks.pop_front();
        }
    };

    HostTimer ht;
    ht.start();
    bcast(0);
    fetch_cols_upto(slab_begin[1]);                       // slab 0 (all of C when nslab == 1)
    for (int64_t k = 1; k <= lookahead && k < kt; ++k)
        bcast(k);
    ht.stop(0);

    const int64_t iterations = kt + extra_panels;
    for (int64_t t = 0; t < iterations; ++t) {
        ht.start();
        if (t > 0 && t + lookahead < kt)
            bcast(t + lookahead);
        fetch_for_iteration(t);
        ht.stop(0);
        ht.start();

        // src/gemmC.cc:113-171: beta only on the first block outer product of a slab
        if (nslab == 1) {
            internal::gemm<Target::Devices>(
                alpha, A.sub(0, A.mt() - 1, t, t), B.sub(t, t, 0, B.nt() - 1),
                t == 0 ? beta : one, std::move(C), layout, 0, queue_0);
        }
        else {
            std::vector<internal::GemmPart<scalar_t>> first, rest;
            for (int64_t s = 0; s < nslab; ++s) {
                const int64_t k = t - skew * s;
                if (k < 0 || k >= kt) continue;
                int64_t j1, j2;
                slab_cols(s, &j1, &j2);
                internal::GemmPart<scalar_t> part{A.sub(0, A.mt() - 1, k, k), B.sub(k, k, j1, j2),
                                                  C.sub(0, C.mt() - 1, j1, j2)};
                (k == 0 ? first : rest).push_back(part);
            }
            internal::gemm_multi(internal::TargetType<Target::Devices>(), alpha, first, beta, layout,
                                 queue_0);
            internal::gemm_multi(internal::TargetType<Target::Devices>(), alpha, rest, one, layout,
                                 queue_0);
            // a slab that has just had its last update goes home once that launch is done
            for (int64_t s = 0; s < nslab; ++s) {
                if (t - skew * s == kt - 1) {
                    int64_t j1, j2;
                    slab_cols(s, &j1, &j2);
                    if (j2 >= j1)
                        writebacks.push_back({j1, j2 + 1, Engine::instance().next_queue_seq(queue_0) - 1});
                }
            }
            drain_writebacks(std::max<int64_t>(2, C.nt() / 16));
        }
        ht.stop(1);
        ht.start();

        // src/gemmC.cc:173-185: drop the panel the last slab has just used (its blocks return
        // to the pool when the update is done)
        const int64_t kr = t - extra_panels;
        if (kr >= 0) {
            auto A_panel = A.sub(0, A.mt() - 1, kr, kr);
            A_panel.releaseRemoteWorkspace();
            A_panel.releaseLocalWorkspace();
            auto B_panel = B.sub(kr, kr, 0, B.nt() - 1);
            B_panel.releaseRemoteWorkspace();
            B_panel.releaseLocalWorkspace();
        }
        Engine::instance().poll_deferred();
        ht.stop(2);
    }

    // src/gemmC.cc:187-190
    ht.start();
    // slabs still waiting to go home: each as soon as its last launch has finished
    while (!writebacks.empty()) {
        drain_writebacks(C.nt());
        if (!writebacks.empty()) std::this_thread::sleep_for(std::chrono::microseconds(100));
    }
    C.tileUpdateAllOrigin();
    C.releaseWorkspace();
    Engine::instance().sync_all();
    ht.stop(3);
    Engine::instance().trace_dump("slate::gemm");
    if (ht.on)
        std::fprintf(stderr, "[slate_b200 rank %d] gemm host seconds: bcast %.3f  internal::gemm %.3f  "
                     "release %.3f  final sync %.3f  (kt = %lld)\n", C.mpiRank(), ht.acc[0], ht.acc[1],
                     ht.acc[2], ht.acc[3], (long long)kt);
}

//------------------------------------------------------------------------------
template <typename scalar_t>
int64_t potrf(HermitianMatrix<scalar_t>& A_in, Options const& opts)
{
    if (get_target(opts) != Target::Devices)
        slate_not_implemented("slate::potrf: this build implements Target::Devices only "
                              "(there is no CPU fallback)");
    using real_t = real_type<scalar_t>;
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    const scalar_t one = 1.0;
    const Layout layout = Layout::ColMajor;
    const int64_t lookahead = get_option<int64_t>(opts, Option::Lookahead, 1);
    const bool hold_local_workspace = get_option<int64_t>(opts, Option::HoldLocalWorkspace, 0) != 0;
    Engine& eng = Engine::instance();

    // trailing update on queue 0; potrf + trsm of the panel on queue 1 (high priority);
    // lookahead column j of step k on queue 2 + (j - k - 1)   (src/potrf.cc:34-36, 152-174)
    const int queue_trail = 0, queue_panel = 1, queue_la0 = 2;
    slate_assert(lookahead >= 0 && queue_la0 + lookahead <= Engine::MaxQueues - 1);

    // if upper, change to lower (src/potrf.cc:45-47)
    HermitianMatrix<scalar_t> A = A_in;
    if (A.uplo() == Uplo::Upper)
        A = conj_transpose(A);
    const int64_t A_nt = A.nt();
    Engine::instance().reserve_info_slots(A_nt);      // one device info per block column

    A.allocateBatchArrays(0, queue_la0 + lookahead);
    A.reserveDeviceWorkspace();
    if (A_nt == 0) return 0;

    // The panel chain (lookahead update, tile potrf, trsm) runs on high-priority queues while
    // the trailing update fills the GPU; each of its kernels starts when a CTA of the update
    // exits.  Shorter-lived update CTAs shorten those waits (tile_engine.h,
    // te_gemm_set_items_per_cta: 117.0 -> 121.2 TFLOP/s on 4 GPUs at n = 98304).
    // On one GPU the chain has no broadcasts in it and hides anyway; there the default cap is
    // the faster one (n = 98304: 33.2 vs 32.7 TFLOP/s), so the hint is for multi-rank runs.
    struct ShortCtaGuard {
        explicit ShortCtaGuard(bool on) : on_(on) { if (on_) te_gemm_set_items_per_cta(4); }
        ~ShortCtaGuard() { if (on_) te_gemm_set_items_per_cta(0); }
        bool on_;
    } short_cta_guard(A.mpiComm()->size > 1);
    static const bool tail_policy_env = [] {
        const char* e = std::getenv("SLATE_B200_POTRF_TAIL_CAP");
        return !(e && e[0] == '0');
    }();
    const bool tail_policy = tail_policy_env && A.mpiComm()->size > 1;

    // info of step k stays on the device (the reference copies it back and syncs each step)
    slate_te_call(te_memset_async(eng.info_slot(0), 0, sizeof(int) * A_nt,
                                  eng.compute_stream(queue_panel)));
    eng.mark_queue(queue_panel);

    for (int64_t k = 0; k < A_nt; ++k) {
        // ---- panel (src/potrf.cc:90-133) ----
        internal::potrf<Target::Devices>(A.sub(k, k), 0, queue_panel, eng.info_slot((int)k));

        if (k + 1 <= A_nt - 1) {
            // send A(k, k) down col A(k+1:nt-1, k)
            BcastListTag bcast_diag = potrf_bcast_list(0, A, k);
            A.template listBcastMT<Target::Devices>(bcast_diag, layout);

            // A(k+1:nt-1, k) * A(k, k)^{-H}
            auto Akk = A.sub(k, k);
            auto Tkk = TriangularMatrix<scalar_t>(Diag::NonUnit, Akk);
            internal::trsm<Target::Devices>(Side::Right, one, conj_transpose(Tkk),
                                            A.sub(k + 1, A_nt - 1, k, k), 0, layout, queue_panel);

            // send A(i, k) across row A(i, k+1:i) and down col A(i:nt-1, i)
            BcastListTag bcast_list_A = potrf_bcast_list(1, A, k);
            A.template listBcastMT<Target::Devices>(bcast_list_A, layout);
        }

        // ---- lookahead columns first: the next panel waits for them (src/potrf.cc:156-177)
        for (int64_t j = k + 1; j < k + 1 + lookahead && j < A_nt; ++j) {
            const int queue_j = queue_la0 + (int)(j - k - 1);
            // A(j, j) -= A(j, k) * A(j, k)^H
            internal::herk<Target::Devices>(real_t(-1.0), A.sub(j, j, k, k), real_t(1.0),
                                            A.sub(j, j), 0, queue_j, layout);
            // A(j+1:nt, j) -= A(j+1:nt-1, k) * A(j, k)^H
            if (j + 1 <= A_nt - 1) {
                auto Ajk = A.sub(j, j, k, k);
                internal::gemm<Target::Devices>(-one, A.sub(j + 1, A_nt - 1, k, k),
                                                conj_transpose(Ajk), one,
                                                A.sub(j + 1, A_nt - 1, j, j), layout, 0, queue_j);
            }
        }

        // ---- trailing submatrix (src/potrf.cc:136-149) ----
        if (k + 1 + lookahead < A_nt) {
            if (tail_policy) {
                // Late in the factorisation the update is shorter than the panel chain that
                // runs beside it: then what matters is how soon the chain's kernels find a free
                // SM, i.e. how short the update's CTAs are, not the update's own L2 locality.
                const int64_t r = A_nt - (k + 1 + lookahead);
                const int64_t local_tiles = r * (r + 1) / 2 / A.mpiComm()->size;
                te_gemm_set_items_per_cta(local_tiles < 300 ? 1 : local_tiles < 800 ? 2 : 4);
            }
            internal::herk<Target::Devices>(real_t(-1.0), A.sub(k + 1 + lookahead, A_nt - 1, k, k),
                                            real_t(1.0), A.sub(k + 1 + lookahead, A_nt - 1), 0,
                                            queue_trail, layout);
        }

        // ---- release the panel (src/potrf.cc:179-192); blocks return when step k is done
        {
            auto panel = A.sub(k, A_nt - 1, k, k);
            panel.releaseRemoteWorkspace();
            if (A.origin() != Target::Devices) {
                // enqueued behind the panel's last writer on the home stream; the host does not
                // wait per step (the final tileUpdateAllOrigin / sync_all completes the copies,
                // and the device blocks go back to the pool only when they have been read)
                panel.tileUpdateAllOriginAsync();
                panel.releaseLocalWorkspace();
            }
        }
        eng.poll_deferred();
    }

    A.tileUpdateAllOrigin();
    if (!hold_local_workspace)
        A.releaseWorkspace();
    eng.sync_all();
    eng.trace_dump("slate::potrf");

    // first failing step -> LAPACK info (src/potrf.cc:104-105, 208)
    std::vector<int> step_info(A_nt, 0);
    slate_te_call(te_memcpy_async(step_info.data(), eng.info_slot(0), sizeof(int) * A_nt,
                                  TE_COPY_D2H, eng.comm_stream()));
    eng.sync_comm();
    int64_t info = 0, kk = 0;
    for (int64_t k = 0; k < A_nt; ++k) {
        if (step_info[k] != 0 && A.tileIsLocal(k, k)) {
            info = kk + step_info[k];
            break;
        }
        kk += A.tileNb(k);
    }
    internal::reduce_info(&info, A.mpiComm());
    return info;
}

//------------------------------------------------------------------------------
namespace {
template <typename scalar_t>
std::vector<int64_t> row_offsets(BaseMatrix<scalar_t>& A)
{
    std::vector<int64_t> off(A.mt() + 1, 0);
    for (int64_t i = 0; i < A.mt(); ++i) off[i + 1] = off[i] + A.tileMb(i);
    return off;
}
template <typename scalar_t>
std::vector<int64_t> col_offsets(BaseMatrix<scalar_t>& A)
{
    std::vector<int64_t> off(A.nt() + 1, 0);
    for (int64_t j = 0; j < A.nt(); ++j) off[j + 1] = off[j] + A.tileNb(j);
    return off;
}
template <typename scalar_t>
bool stored(BaseMatrix<scalar_t>& A, int64_t i, int64_t j)
{
    if (A.uploPhysical() == Uplo::General) return true;
    return A.uploLogical() == Uplo::Lower ? i >= j : i <= j;
}
}  // namespace

template <typename scalar_t>
void generate_rand(BaseMatrix<scalar_t>& A, int64_t seed, bool dominant, int dist)
{
    slate_assert(A.op() == Op::NoTrans);
    Engine& eng = Engine::instance();
    const int dtype = internal::te_dtype_of<scalar_t>::value;
    auto roff = row_offsets(A);
    auto coff = col_offsets(A);
    const double diag_add = dominant ? (double)roff[A.mt()] : 0.0;
    void* stream = eng.compute_stream(0);
    void* staging = nullptr;
    for (int64_t j = 0; j < A.nt(); ++j) {
        for (int64_t i = 0; i < A.mt(); ++i) {
            if (!A.tileIsLocal(i, j) || !stored(A, i, j)) continue;
            const int mb = (int)A.tileMb(i), nb = (int)A.tileNb(j);
            if (A.tileExists(i, j, 0) && A.origin() == Target::Devices) {
                auto T = A(i, j, 0);
                slate_te_call(te_matgen_tile(eng.device(), stream, dtype, dist, seed, mb, nb, roff[i],
                                             coff[j], T.data(), T.stride(), diag_add));
                A.tileModified(i, j, 0, 0);
            }
            else {
                // host-origin tile: same bits, produced on the device and copied back
                if (!staging)
                    staging = eng.scratch(Engine::MaxQueues - 2,
                                          sizeof(scalar_t) * (size_t)A.storage()->mb_ * A.storage()->nb_);
                auto T = A(i, j, HostNum);
                slate_te_call(te_matgen_tile(eng.device(), stream, dtype, dist, seed, mb, nb, roff[i],
                                             coff[j], staging, mb, diag_add));
                slate_te_call(te_memcpy2d_async(T.data(), sizeof(scalar_t) * T.stride(), staging,
                                                sizeof(scalar_t) * mb, sizeof(scalar_t) * mb, nb,
                                                TE_COPY_D2H, stream));
                eng.sync_queue(0);
                A.tileModified(i, j, HostNum);
            }
        }
    }
    eng.mark_queue(0);
    eng.sync_queue(0);
}

template <typename scalar_t>
double local_sumsq(BaseMatrix<scalar_t>& A)
{
    Engine& eng = Engine::instance();
    const int dtype = internal::te_dtype_of<scalar_t>::value;
    void* stream = eng.compute_stream(0);
    double* acc = (double*)eng.scratch(Engine::MaxQueues - 3, sizeof(double));
    slate_te_call(te_memset_async(acc, 0, sizeof(double), stream));
    const bool trapezoid = A.uploPhysical() != Uplo::General;
    for (int64_t j = 0; j < A.nt(); ++j) {
        for (int64_t i = 0; i < A.mt(); ++i) {
            if (!A.tileIsLocal(i, j) || !stored(A, i, j)) continue;
            A.tileGetForReading(i, j, 0, LayoutConvert::None, 0);
            Engine::instance().mark_comm();
            Engine::instance().queue_wait_comm(0);
            auto T = A(i, j, 0);
            T.op(Op::NoTrans);
            const int mb = (int)T.mb(), nb = (int)T.nb();
            if (!trapezoid) {
                slate_te_call(te_tile_sumsq(eng.device(), stream, dtype, 0, mb, nb, T.data(),
                                            T.stride(), acc));
            }
            else if (i == j) {
                // full Hermitian norm: diagonal once, strict triangle twice
                const bool lower = A.uploPhysical() == Uplo::Lower;
                slate_te_call(te_tile_sumsq(eng.device(), stream, dtype, lower ? 1 : 2, mb, nb,
                                            T.data(), T.stride(), acc));
                if (lower)
                    slate_te_call(te_tile_sumsq(eng.device(), stream, dtype, 3, mb, nb, T.data(),
                                                T.stride(), acc));
                else
                    slate_error("local_sumsq: upper storage not supported");
            }
            else {
                slate_te_call(te_tile_sumsq(eng.device(), stream, dtype, 0, mb, nb, T.data(),
                                            T.stride(), acc));
                slate_te_call(te_tile_sumsq(eng.device(), stream, dtype, 0, mb, nb, T.data(),
                                            T.stride(), acc));
            }
        }
    }
    double result = 0;
    slate_te_call(te_memcpy_async(&result, acc, sizeof(double), TE_COPY_D2H, stream));
    eng.mark_queue(0);
    eng.sync_queue(0);
    return result;
}

template <typename scalar_t>
double potrf_residual_sumsq(HermitianMatrix<scalar_t>& A, int64_t seed, bool dominant,
                            double* a0_sumsq)
{
    using real_t = real_type<scalar_t>;
    using BcastListTag = typename Matrix<scalar_t>::BcastListTag;
    slate_assert(A.uploPhysical() == Uplo::Lower && A.op() == Op::NoTrans);
    slate_assert(A.origin() == Target::Devices);
    Engine& eng = Engine::instance();
    const int dtype = internal::te_dtype_of<scalar_t>::value;
    auto st = A.storage();
    const int64_t n = st->n_, nb = st->nb_;
    const int64_t nt = A.nt();
    const Layout layout = Layout::ColMajor;

    // R = A0, regenerated bit for bit
    HermitianMatrix<scalar_t> R(Uplo::Lower, n, nb, st->p_, st->q_, st->comm_);
    R.insertLocalTiles(Target::Devices);
    R.reserveDeviceWorkspace();
    generate_rand(R, seed, dominant, 1);
    if (a0_sumsq) *a0_sumsq = local_sumsq(R);

    // shadow matrix holding clean copies of block column k of L (strict upper of L(k,k) zeroed)
    Matrix<scalar_t> Lm(n, n, nb, st->p_, st->q_, st->comm_);
    Lm.reserveDeviceWorkspace(3 * nt + 8);
    void* stream = eng.compute_stream(0);
    for (int64_t k = 0; k < nt; ++k) {
        for (int64_t i = k; i < nt; ++i) {
            if (!A.tileIsLocal(i, k)) continue;
            A.tileGetForReading(i, k, 0, LayoutConvert::None, 0);
            auto src = A(i, k, 0);
            Lm.tileInsert(i, k, 0);
            auto dst = Lm(i, k, 0);
            slate_assert(src.stride() == src.mb() && dst.stride() == dst.mb());
            slate_te_call(te_memcpy_async(dst.data(), src.data(), src.bytes(), TE_COPY_D2D, stream));
            if (i == k)
                slate_te_call(te_tile_keep_triangle(eng.device(), stream, dtype, TE_UPLO_LOWER,
                                                    (int)dst.mb(), (int)dst.nb(), dst.data(),
                                                    dst.stride()));
            Lm.tileModified(i, k, 0, 0);
        }
        eng.mark_queue(0);
        BcastListTag list;
        for (int64_t i = k; i < nt; ++i)
            list.push_back({i, k, {R.sub(i, i, k, i), R.sub(i, nt - 1, i, i)}, i});
        Lm.template listBcastMT<Target::Devices>(list, layout);
        // R(k:, k:) -= L(k:, k) L(k:, k)^H
        internal::herk<Target::Devices>(real_t(-1.0), Lm.sub(k, nt - 1, k, k), real_t(1.0),
                                        R.sub(k, nt - 1), 0, 0, layout);
        auto panel = Lm.sub(k, nt - 1, k, k);
        panel.releaseRemoteWorkspace();
        for (int64_t i = k; i < nt; ++i)
            if (Lm.tileIsLocal(i, k)) Lm.tileErase(i, k, 0);
        Lm.storage()->flushErased();
        eng.poll_deferred();
    }
    double r = local_sumsq(R);
    eng.sync_all();
    return r;
}

//------------------------------------------------------------------------------
#define INSTANTIATE(T)                                                                        \
    template typename Matrix<T>::BcastListTag gemm_bcast_list<T>(int, Matrix<T>&, Matrix<T>&,  \
                                                                 Matrix<T>&, int64_t);         \
    template typename Matrix<T>::BcastListTag potrf_bcast_list<T>(int, HermitianMatrix<T>&,    \
                                                                  int64_t);                    \
    template void gemmC<T>(T, Matrix<T>&, Matrix<T>&, T, Matrix<T>&, Options const&);          \
    template int64_t potrf<T>(HermitianMatrix<T>&, Options const&);                            \
    template void generate_rand<T>(BaseMatrix<T>&, int64_t, bool, int);                        \
    template double local_sumsq<T>(BaseMatrix<T>&);                                            \
    template double potrf_residual_sumsq<T>(HermitianMatrix<T>&, int64_t, bool, double*);
INSTANTIATE(double)
INSTANTIATE(std::complex<double>)

}  // namespace slate
