// pj-learn -- drop-in for the reference tool (src/pj-learn.cpp): nuclear-norm regularised RDA over
// the unprojected distances.  Same argv, same outputs (`W` rank x D, `A` D x D, contiguous f32) and
// stdout lines.  Batch sampling (:311-329) stays host code (cv::RNG stream); margins, hinge mask,
// the weighted SYRK, the dual-average update, the eigen-decomposition (cuSOLVER) and the prox run
// in libdlco_b200.so, as do the validation objective (:501-527) and ComputePJStats.
//
//   pj-learn src_h5_dist_file dst_h5_output_file [-mu m] [-gamma g] [-iters n] [-prox eig|sign]
//
// -prox sign (or DLCO_PJ_PROX=sign in the environment, so workspace/08-pjlearn.sh needs no edit): the prox's
// projection by the tensor-core matrix-sign iteration instead of an eigen-decomposition per iteration
// (dlco_pj_rda_step_grid_sign; sampling on the device with the same cv::RNG stream); W is rebuilt from the
// projected A at the validation points.  ~10x more iterations per second, same files and log lines.
#include "common.hpp"

#include <cfloat>

using namespace dlco;

int main(int argc, char **argv)
{
    float mu = 0.001f, gamma = 0.5f;
    unsigned int n_iter = 50000;
    const unsigned int log_step = 100, sz_batch = 200;
    bool help = false;
    const char *prox_env = getenv("DLCO_PJ_PROX");
    bool prox_sign = prox_env && !strcmp(prox_env, "sign");
    const char *dists = nullptr, *outp = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-mu")) { mu = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gamma")) { gamma = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-iters")) { n_iter = (unsigned int)atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prox")) { prox_sign = !strcmp(argv[++i], "sign"); continue; }
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
