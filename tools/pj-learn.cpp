// pj-learn -- drop-in for the reference tool (src/pj-learn.cpp): nuclear-norm regularised RDA over
// the unprojected distances.  Same argv, same outputs (`W` rank x D, `A` D x D, contiguous f32) and
// stdout lines.  Batch sampling (:311-329) stays host code (cv::RNG stream); margins, hinge mask,
// the weighted SYRK, the dual-average update, the eigen-decomposition (cuSOLVER) and the prox run
// in libdlco_b200.so, as do the validation objective (:501-527) and ComputePJStats.
//
//   pj-learn src_h5_dist_file dst_h5_output_file [-mu m] [-gamma g] [-iters n] [-prox eig|sign] [-gpus N]
//
// -gpus N (extension): the pairs ("Distance" rows) are sharded over N GPUs of the box; every iteration the margins of
// the sampled rows and the D x D gradient are all-reduced over NVLink peer memory (dlco_comm_allreduce_f32), the update
// and the prox run on every GPU on identical data.  Same files and log lines as one GPU.
//
// -prox sign (or DLCO_PJ_PROX=sign in the environment, so workspace/08-pjlearn.sh needs no edit): the prox's
// projection by the tensor-core matrix-sign iteration instead of an eigen-decomposition per iteration
// (dlco_pj_rda_step_grid_sign; sampling on the device with the same cv::RNG stream); W is rebuilt from the
// projected A at the validation points.  ~10x more iterations per second, same files and log lines.
#include "common.hpp"

#include <cfloat>
#include <memory>

using namespace dlco;

// [lo, hi) of rank g: contiguous, boundaries on multiples of 128 rows (the HDF5 chunk), as opencv-dlco_b200/shard.py
static void pair_range(uint64_t n, int g, int world, uint64_t *lo, uint64_t *hi)
{
    const uint64_t blocks = (n + 127) / 128, per = blocks / world, extra = blocks % world;
    const uint64_t lo_b = (uint64_t)g * per + std::min<uint64_t>(g, extra), hi_b = lo_b + per + ((uint64_t)g < extra ? 1 : 0);
    *lo = std::min(n, lo_b * 128);
    *hi = std::min(n, hi_b * 128);
}

// One process, N GPUs, pairs sharded (north_star: "sharding training pairs, with an all-reduce over NVLink of the per-step
// gradient").  Per iteration and GPU: margins of the sampled rows it owns -> all-reduce of the 2 x szBatch margins ->
// hinge weights -> weighted SYRK over the owned rows -> all-reduce of the D x D gradient -> update, eig, prox (redundant,
// bit-identical on every GPU).  Validation and statistics: margins / patch distances of the owned rows, all-reduced into
// full-length vectors, then GPU 0 alone.  A single host thread queues the work of all GPUs; the all-reduce kernels wait
// for each other on the devices.
static int run_pair_sharded(const char *dists, const char *outp, float mu, float gamma, unsigned int n_iter,
                            unsigned int log_step, unsigned int sz_batch, int n_gpus)
{
    try {
        const int io_threads = host_threads();
        h5mini::Reader r(dists);
        const h5mini::DatasetInfo &di = r.info("Distance");
        if (di.dtype != h5mini::F32 || di.dims.size() != 2) throw h5mini::Error("Distance must be a 2-D f32 dataset");
        const uint64_t N = di.dims[0];
        const int D = (int)di.dims[1];
        printf("Load Labels: %llu\nLoad Distances: %llu x %llu\n", (unsigned long long)N, (unsigned long long)N, (unsigned long long)D);
        std::vector<uint8_t> labels(N);
        if (N) r.read_rows("Label", 0, N, labels.data(), io_threads);
        struct Gpu {
            std::unique_ptr<Ctx> ctx;
            uint64_t lo = 0, hi = 0;
            DevBuf<float> dists, W, A, dloss, part, dfavg, evec, eval, m_loc, m_full, c, c_loc, pd_loc, vec;
            DevBuf<int32_t> pos, loc, own_rows;
            DevBuf<uint8_t> labels;
            dlco_comm *comm = nullptr;
        };
        std::vector<Gpu> G((size_t)n_gpus);
        std::vector<dlco_ctx *> ctxs;
        const size_t DD = (size_t)D * D;
        Progress bar;
        for (int g = 0; g < n_gpus; g++) {
            Gpu &u = G[g];
            u.ctx.reset(new Ctx(g));
            ctxs.push_back(u.ctx->h);
            pair_range(N, g, n_gpus, &u.lo, &u.hi);
            const uint64_t rows = u.hi - u.lo;
            u.dists.alloc(*u.ctx, std::max<uint64_t>(rows, 1) * D);
            std::vector<float> buf;
            const uint64_t blk = 16384;
            for (uint64_t i = u.lo; i < u.hi; i += blk) {
                const uint64_t cnt = std::min(blk, u.hi - i);
                buf.resize(cnt * D);
                r.read_rows("Distance", i, i + cnt, buf.data(), io_threads);
                u.dists.upload(buf.data(), cnt * D, (i - u.lo) * D);
                bar.update((double)i / (double)N);
            }
            for (DevBuf<float> *b : {&u.W, &u.A, &u.dloss, &u.part, &u.dfavg, &u.evec}) b->alloc(*u.ctx, DD);
            u.eval.alloc(*u.ctx, D);
            std::vector<float> z(DD, 0.0f);
            u.W.upload(z.data(), DD);
            u.A.upload(z.data(), DD);
            u.dfavg.upload(z.data(), DD);
            u.m_loc.alloc(*u.ctx, std::max<uint64_t>(rows, 2 * sz_batch));
            u.m_full.alloc(*u.ctx, 2 * sz_batch);
            u.c.alloc(*u.ctx, 2 * sz_batch);
            u.c_loc.alloc(*u.ctx, 2 * sz_batch);
            u.pos.alloc(*u.ctx, std::max<uint64_t>(N, 2 * sz_batch));
            u.loc.alloc(*u.ctx, std::max<uint64_t>(N, 2 * sz_batch));
            u.vec.alloc(*u.ctx, std::max<uint64_t>(N, 1));
            u.pd_loc.alloc(*u.ctx, std::max<uint64_t>(rows, 1));
        }
        bar.update(1.0);
        Split sp = split_train_val(labels);
        printf("Positive samples #%zu\nNegative samples #%zu\n", sp.idx_pos.size(), sp.idx_neg.size());
        printf("Positive train #%zu\nNegative train #%zu\nPositive valid #%zu\nNegative valid #%zu\n", sp.n_pos_trn,
               sp.n_neg_trn, sp.n_pos_val(), sp.n_neg_val());
        if (!sp.n_pos_trn || !sp.n_neg_trn) throw h5mini::Error("need positive and negative training pairs");
        G[0].ctx->print_gpu();
        std::vector<dlco_comm *> comms((size_t)n_gpus, nullptr);
        check(dlco_comm_create_local(ctxs.data(), n_gpus, std::max<size_t>(DD, N), comms.data()), "dlco_comm_create_local");
        for (int g = 0; g < n_gpus; g++) G[g].comm = comms[g];
        fprintf(stderr, "[pj-learn] %d GPUs, pairs sharded in ranges of %llu rows, all-reduce over peer memory\n", n_gpus,
                (unsigned long long)(G[0].hi - G[0].lo));
        G[0].labels.alloc(*G[0].ctx, N);
        G[0].labels.upload(labels.data(), N);

        // full-length vector of a per-row quantity over an explicit global row list: every GPU fills the entries of the
        // rows it owns, the all-reduce completes it everywhere (each slot has exactly one owner)
        std::vector<int32_t> h_pos, h_loc;
        auto rows_quantity = [&](const std::vector<int32_t> &rows, int rank_w, bool queue_only) {
            for (int g = 0; g < n_gpus; g++) {
                Gpu &u = G[g];
                check(dlco_make_current(u.ctx->h), "dlco_make_current");
                h_pos.clear();
                h_loc.clear();
                for (size_t k = 0; k < rows.size(); k++)
                    if ((uint64_t)rows[k] >= u.lo && (uint64_t)rows[k] < u.hi) {
                        h_pos.push_back((int32_t)k);
                        h_loc.push_back((int32_t)(rows[k] - u.lo));
                    }
                const int n_own = (int)h_pos.size();
                u.pos.upload(h_pos.data(), h_pos.size());
                u.loc.upload(h_loc.data(), h_loc.size());
                if (n_own)
                    check(dlco_pj_margins(u.ctx->h, u.dists.p, D, D, u.loc.p, n_own, u.W.p, rank_w, u.m_loc.p), "dlco_pj_margins");
                check(dlco_scatter_f32(u.ctx->h, u.m_loc.p, u.pos.p, n_own, u.vec.p, (int)rows.size()), "dlco_scatter_f32");
                check(dlco_comm_allreduce_f32(u.comm, u.vec.p, u.vec.p, rows.size()), "dlco_comm_allreduce_f32");
                (void)queue_only;
            }
        };

        std::vector<int32_t> rows(2 * sz_batch), val_rows;
        for (size_t k = sp.n_pos_trn; k < sp.idx_pos.size(); k++) val_rows.push_back(sp.idx_pos[k]);
        for (size_t k = sp.n_neg_trn; k < sp.idx_neg.size(); k++) val_rows.push_back(sp.idx_neg[k]);
        std::vector<float> W_save, A_save;
        int rank_save = -1;
        CvRng rng(2215);
        int rank = 0;
        float obj_best = FLT_MAX, fpr95_best = 1.0f;
        double auc_best = 0.0;
        int rows_best = 0;
        unsigned int step = 0;
        double t_train = now_sec();
        std::vector<int> n_own((size_t)n_gpus);
        for (unsigned int t = 0; t <= n_iter; t++) {
            for (unsigned int k = 0; k < sz_batch; k++) {        // interleaved pos, neg draws (:311-316)
                rows[k] = sp.idx_pos[(size_t)rng.uniform(0, (int)sp.n_pos_trn)];
                rows[sz_batch + k] = sp.idx_neg[(size_t)rng.uniform(0, (int)sp.n_neg_trn)];
            }
            // margins of the owned sampled rows -> all-reduce of the 2 x szBatch margins
            for (int g = 0; g < n_gpus; g++) {
                Gpu &u = G[g];
                check(dlco_make_current(u.ctx->h), "dlco_make_current");
                h_pos.clear();
                h_loc.clear();
                for (unsigned int k = 0; k < 2 * sz_batch; k++)
                    if ((uint64_t)rows[k] >= u.lo && (uint64_t)rows[k] < u.hi) {
                        h_pos.push_back((int32_t)k);
                        h_loc.push_back((int32_t)(rows[k] - u.lo));
                    }
                n_own[g] = (int)h_pos.size();
                u.pos.upload(h_pos.data(), h_pos.size());
                u.loc.upload(h_loc.data(), h_loc.size());
                if (n_own[g])
                    check(dlco_pj_margins(u.ctx->h, u.dists.p, D, D, u.loc.p, n_own[g], u.W.p, rank, u.m_loc.p), "dlco_pj_margins");
                check(dlco_scatter_f32(u.ctx->h, u.m_loc.p, u.pos.p, n_own[g], u.m_full.p, (int)(2 * sz_batch)), "dlco_scatter_f32");
                check(dlco_comm_allreduce_f32(u.comm, u.m_full.p, u.m_full.p, 2 * sz_batch), "dlco_comm_allreduce_f32");
            }
            // hinge weights (redundant) -> partial gradient over the owned rows -> all-reduce of the D x D matrix
            for (int g = 0; g < n_gpus; g++) {
                Gpu &u = G[g];
                check(dlco_make_current(u.ctx->h), "dlco_make_current");
                check(dlco_pj_hinge_weights(u.ctx->h, u.m_full.p, (int)sz_batch, u.c.p), "dlco_pj_hinge_weights");
                check(dlco_gather_f32(u.ctx->h, u.c.p, u.pos.p, n_own[g], u.c_loc.p), "dlco_gather_f32");
                check(dlco_pj_wsyrk(u.ctx->h, u.dists.p, D, D, u.loc.p, n_own[g], u.c_loc.p, u.part.p), "dlco_pj_wsyrk");
                check(dlco_comm_allreduce_f32(u.comm, u.part.p, u.dloss.p, DD), "dlco_comm_allreduce_f32");
            }
            // dual average, eigen-decomposition, prox: on every GPU, on identical data
            for (int g = 0; g < n_gpus; g++) {
                Gpu &u = G[g];
                check(dlco_make_current(u.ctx->h), "dlco_make_current");
                check(dlco_pj_rda_update(u.ctx->h, D, t, (int)sz_batch, mu, gamma, u.dloss.p, u.dfavg.p, u.A.p), "dlco_pj_rda_update");
                check(dlco_pj_eig(u.ctx->h, D, u.A.p, u.eval.p, u.evec.p), "dlco_pj_eig");
            }
            for (int g = 0; g < n_gpus; g++) {
                Gpu &u = G[g];
                int rk = 0;
                check(dlco_make_current(u.ctx->h), "dlco_make_current");
                check(dlco_pj_prox_apply(u.ctx->h, D, u.eval.p, u.evec.p, u.A.p, u.W.p, &rk), "dlco_pj_prox_apply");
                if (g == 0) rank = rk;
                else if (rk != rank) throw AbiError("the GPUs disagree on the rank of W: the replicas have diverged");
            }
            if (step == log_step) {
                for (int g = 0; g < n_gpus; g++) check(dlco_comm_check(G[g].comm), "dlco_comm_check");
                const double t_valid = now_sec();
                // validation objective (:501-527): margins of the hold-out rows, completed on every GPU, hinge on GPU 0
                rows_quantity(val_rows, rank, false);
                Gpu &u0 = G[0];
                check(dlco_make_current(u0.ctx->h), "dlco_make_current");
                double hs = 0.0;
                check(dlco_hinge_pairsum(u0.ctx->h, u0.vec.p, (int64_t)sp.n_pos_val(), u0.vec.p + sp.n_pos_val(), (int64_t)sp.n_neg_val(),
                                         1.0f, &hs), "dlco_hinge_pairsum");
                const float loss = (float)hs / (float)sp.n_pos_val() / (float)sp.n_neg_val();
                std::vector<float> A_host(DD);
                u0.A.download(A_host.data(), DD);
                double tr = 0.0;
                for (int i = 0; i < D; i++) tr += (double)A_host[(size_t)i * D + i];
                const float regul = (float)((double)mu * tr);
                const double t_done = now_sec();
                const float obj = loss + regul;
                const int w_rows = rank > 0 ? rank : D;
                if (obj < obj_best) {
                    obj_best = obj;
                    rows_best = w_rows;
                    printf("Best: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t, loss,
                           regul, obj, obj_best, w_rows, rows_best, t_valid - t_train, t_done - t_valid);
                    // ComputePJStats distances of every pair: each GPU's rows are one contiguous slice of the vector
                    for (int g = 0; g < n_gpus; g++) {
                        Gpu &u = G[g];
                        check(dlco_make_current(u.ctx->h), "dlco_make_current");
                        check(dlco_scatter_f32(u.ctx->h, nullptr, nullptr, 0, u.vec.p, (int)N), "dlco_scatter_f32");     // zero
                        if (u.hi > u.lo)
                            check(dlco_pj_patchdist(u.ctx->h, u.dists.p, (int64_t)(u.hi - u.lo), D, D, u.W.p, rank, u.vec.p + u.lo),
                                  "dlco_pj_patchdist");
                        check(dlco_comm_allreduce_f32(u.comm, u.vec.p, u.vec.p, N), "dlco_comm_allreduce_f32");
                    }
                    check(dlco_make_current(u0.ctx->h), "dlco_make_current");
                    float fpr95 = 0.f;
                    double auc = 0.0;
                    check(dlco_roc_stats(u0.ctx->h, u0.vec.p, u0.labels.p, (int64_t)N, &fpr95, &auc), "dlco_roc_stats");
                    if (auc_best <= auc && fpr95_best >= fpr95) {
                        auc_best = auc;
                        fpr95_best = fpr95;
                        rank_save = w_rows;
                        W_save.resize((size_t)w_rows * D);
                        u0.W.download(W_save.data(), W_save.size());
                        A_save = A_host;
                        printf("Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f) [saved]\n", rank, auc, auc_best, fpr95 * 100,
                               fpr95_best * 100);
                    } else {
                        printf("Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f)\n", rank, auc, auc_best, fpr95 * 100,
                               fpr95_best * 100);
                    }
                } else {
                    printf("Step: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t, loss,
                           regul, obj, obj_best, w_rows, rows_best, t_valid - t_train, t_done - t_valid);
                }
                fflush(stdout);
                step = 0;
                t_train = now_sec();
            }
            step++;
        }
        for (int g = 0; g < n_gpus; g++) {
            check(dlco_comm_check(G[g].comm), "dlco_comm_check");
            dlco_comm_destroy(G[g].comm);
        }
        if (rank_save < 0) {
            rank_save = D;
            W_save.assign(DD, 0.0f);
            A_save.assign(DD, 0.0f);
        }
        h5mini::Writer out(outp);
        out.write_contiguous("W", h5mini::F32, {(uint64_t)rank_save, (uint64_t)D}, W_save.data());
        out.write_contiguous("A", h5mini::F32, {(uint64_t)D, (uint64_t)D}, A_save.data());
        out.close();
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npj-learn: %s\n", e.what());
        return 2;
    }
}

int main(int argc, char **argv)
{
    float mu = 0.001f, gamma = 0.5f;
    unsigned int n_iter = 50000;
    const unsigned int log_step = 100, sz_batch = 200;
    bool help = false;
    const char *prox_env = getenv("DLCO_PJ_PROX");
    bool prox_sign = prox_env && !strcmp(prox_env, "sign");
    int n_gpus = 1;
    const char *dists = nullptr, *outp = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-mu")) { mu = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gamma")) { gamma = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-iters")) { n_iter = (unsigned int)atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prox")) { prox_sign = !strcmp(argv[++i], "sign"); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gpus")) { n_gpus = std::max(1, atoi(argv[++i])); continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!dists) dists = argv[i];
        else if (!outp) outp = argv[i];
    }
    if (!dists || !outp) help = true;
    if (help) {
        // the reference prints "pr-learn" and pr-learn's defaults here (src/pj-learn.cpp:159-163)
        printf("\nUsage: pr-learn  src_h5_dist_file dst_h5_output_file\n"
               "       -mu <0.0-1.0, 0.025=default> \n"
               "       -gamma <0.0-10.0, 0.10=default> \n"
               "       -iters <0-N, 5000000=default> \n\n");
        return 1;
    }
    printf("mu: %g gamma: %g nIters: %u\n", mu, gamma, n_iter);
    n_iter = max_iters_cap(n_iter);                        // DLCO_MAX_ITERS, see tools/common.hpp
    if (n_gpus > 1) return run_pair_sharded(dists, outp, mu, gamma, n_iter, log_step, sz_batch, n_gpus);
    try {
        Ctx ctx(0);
        DevBuf<float> d_dists;
        DistFile df = load_dist_file(dists, ctx, d_dists, true);
        const int D = (int)df.f;
        Split sp = split_train_val(df.labels);
        printf("Positive samples #%zu\nNegative samples #%zu\n", sp.idx_pos.size(), sp.idx_neg.size());
        printf("Positive train #%zu\nNegative train #%zu\nPositive valid #%zu\nNegative valid #%zu\n", sp.n_pos_trn,
               sp.n_neg_trn, sp.n_pos_val(), sp.n_neg_val());
        if (!sp.n_pos_trn || !sp.n_neg_trn) throw h5mini::Error("need positive and negative training pairs");
        ctx.print_gpu();

        DevBuf<int32_t> d_pos(ctx, sp.idx_pos.size()), d_neg(ctx, sp.idx_neg.size());
        d_pos.upload(sp.idx_pos.data(), sp.idx_pos.size());
        d_neg.upload(sp.idx_neg.data(), sp.idx_neg.size());
        DevBuf<uint8_t> d_labels(ctx, df.n);
        d_labels.upload(df.labels.data(), df.n);
        const size_t DD = (size_t)D * D;
        DevBuf<float> d_W(ctx, DD), d_A(ctx, DD), d_dloss(ctx, DD), d_dfavg(ctx, DD), d_evec(ctx, DD), d_eval(ctx, D);
        DevBuf<float> d_pd(ctx, df.n);
        DevBuf<int32_t> d_rows(ctx, 2 * sz_batch);
        {
            std::vector<float> z(DD, 0.0f);
            d_W.upload(z.data(), DD);
            d_A.upload(z.data(), DD);
            d_dfavg.upload(z.data(), DD);
        }
        // -prox sign: device-resident sampler state, (mu, gamma) and rank of the one configuration
        DevBuf<uint64_t> d_rng(ctx, 1);
        DevBuf<float> d_mu(ctx, 1), d_gamma(ctx, 1), d_work(ctx, prox_sign ? DD + D : 1);
        DevBuf<int32_t> d_rank(ctx, 1);
        if (prox_sign) {
            const uint64_t seed = 2215;
            const int32_t r0 = 0;
            d_rng.upload(&seed, 1);
            d_mu.upload(&mu, 1);
            d_gamma.upload(&gamma, 1);
            d_rank.upload(&r0, 1);
        }
        std::vector<int32_t> rows(2 * sz_batch);
        std::vector<float> W_save, A_save;
        int rank_save = -1;

        CvRng rng(2215);                                     // RNG rng( 2215 ), :225
        int rank = 0;                                        // W = zeros(D, D): all margins 0 at t = 0
        float obj_best = FLT_MAX, fpr95_best = 1.0f;
        double auc_best = 0.0;
        int rows_best = 0;
        unsigned int step = 0;
        double t_train = now_sec();
        for (unsigned int t = 0; t <= n_iter; t++) {
            if (prox_sign) {
                check(dlco_pj_rda_step_grid_sign(ctx.h, d_dists.p, D, D, d_pos.p, (int32_t)sp.n_pos_trn, d_neg.p,
                                                 (int32_t)sp.n_neg_trn, (int)sz_batch, d_rng.p, 1, d_mu.p, d_gamma.p, t, d_A.p,
                                                 d_dfavg.p, d_rank.p, d_dloss.p),
                      "dlco_pj_rda_step_grid_sign");
            } else {
                for (unsigned int k = 0; k < sz_batch; k++) {    // interleaved pos, neg draws (:311-316)
                    rows[k] = sp.idx_pos[(size_t)rng.uniform(0, (int)sp.n_pos_trn)];
                    rows[sz_batch + k] = sp.idx_neg[(size_t)rng.uniform(0, (int)sp.n_neg_trn)];
                }
                d_rows.upload(rows.data(), rows.size());
                check(dlco_pj_rda_grad(ctx.h, d_dists.p, D, D, d_rows.p, d_rows.p + sz_batch, (int)sz_batch, d_W.p, rank, d_dloss.p),
                      "dlco_pj_rda_grad");
                check(dlco_pj_rda_update(ctx.h, D, t, (int)sz_batch, mu, gamma, d_dloss.p, d_dfavg.p, d_A.p), "dlco_pj_rda_update");
                check(dlco_pj_eig(ctx.h, D, d_A.p, d_eval.p, d_evec.p), "dlco_pj_eig");
                check(dlco_pj_prox_apply(ctx.h, D, d_eval.p, d_evec.p, d_A.p, d_W.p, &rank), "dlco_pj_prox_apply");
            }

            if (step == log_step) {
                if (prox_sign) {          // W of the projected A for validation, statistics and the saved model
                    check(dlco_pj_factor_grid(ctx.h, D, 1, d_A.p, d_rank.p, d_W.p, (int64_t)DD, d_work.p), "dlco_pj_factor_grid");
                    int32_t r1 = 0;
                    d_rank.download(&r1, 1);
                    rank = r1;
                }
                ctx.sync();
                const double t_valid = now_sec();
                float loss = 0.f, regul = 0.f;
                check(dlco_pj_validate(ctx.h, d_dists.p, D, D, d_pos.p + sp.n_pos_trn, (int64_t)sp.n_pos_val(),
                                       d_neg.p + sp.n_neg_trn, (int64_t)sp.n_neg_val(), d_W.p, rank, d_A.p, mu, &loss, &regul),
                      "dlco_pj_validate");
                const double t_done = now_sec();
                const float obj = loss + regul;
                const int w_rows = rank > 0 ? rank : D;      // an empty W becomes a D x D zero matrix (:487-490)
                if (obj < obj_best) {
                    obj_best = obj;
                    rows_best = w_rows;
                    printf("Best: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t, loss,
                           regul, obj, obj_best, w_rows, rows_best, t_valid - t_train, t_done - t_valid);
                    float fpr95 = 0.f;
                    double auc = 0.0;
                    check(dlco_pj_patchdist(ctx.h, d_dists.p, (int64_t)df.n, D, D, d_W.p, rank, d_pd.p), "dlco_pj_patchdist");
                    check(dlco_roc_stats(ctx.h, d_pd.p, d_labels.p, (int64_t)df.n, &fpr95, &auc), "dlco_roc_stats");
                    if (auc_best <= auc && fpr95_best >= fpr95) {                 // :560-566
                        auc_best = auc;
                        fpr95_best = fpr95;
                        rank_save = w_rows;
                        W_save.resize((size_t)w_rows * D);
                        A_save.resize(DD);
                        d_W.download(W_save.data(), W_save.size());
                        d_A.download(A_save.data(), DD);
                        printf("Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f) [saved]\n", rank, auc, auc_best, fpr95 * 100,
                               fpr95_best * 100);
                    } else {
                        printf("Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f)\n", rank, auc, auc_best, fpr95 * 100,
                               fpr95_best * 100);
                    }
                } else {
                    printf("Step: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t, loss,
                           regul, obj, obj_best, w_rows, rows_best, t_valid - t_train, t_done - t_valid);
                }
                fflush(stdout);
                step = 0;
                t_train = now_sec();
            }
            step++;
        }
        // W_Save / A_Save (src/pj-learn.cpp:592-594); nothing accepted -> the all-zero initial state
        if (rank_save < 0) {
            rank_save = D;
            W_save.assign(DD, 0.0f);
            A_save.assign(DD, 0.0f);
        }
        h5mini::Writer out(outp);
        out.write_contiguous("W", h5mini::F32, {(uint64_t)rank_save, (uint64_t)D}, W_save.data());
        out.write_contiguous("A", h5mini::F32, {(uint64_t)D, (uint64_t)D}, A_save.data());
        out.close();
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npj-learn: %s\n", e.what());
        return 2;
    }
}
