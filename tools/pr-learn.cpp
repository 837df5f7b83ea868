// pr-learn -- drop-in for the reference tool (src/pr-learn.cpp): l1-regularised non-negative RDA
// over the full-distance matrix.  Same argv, HDF5 layout (`w`: unlimited rows x F, one row appended
// per accepted model) and stdout lines; the chain (:303-329), the validation objective (:340-358)
// and ComputePRStats' distance pass + ROC run in libdlco_b200.so.
//
//   pr-learn src_h5_filter_file src_h5_dist_file dst_h5_output_file [-mu m] [-gamma g] [-maxdim d] [-iters n]
#include "common.hpp"

#include <cfloat>
#include <memory>

using namespace dlco;

int main(int argc, char **argv)
{
    float mu = 0.025f, gamma = 0.10f;
    int max_dim = 640;
    unsigned int n_iter = 5000000;
    const unsigned int log_step = 100000;
    bool help = false;
    const char *flt = nullptr, *dists = nullptr, *outp = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-mu")) { mu = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gamma")) { gamma = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-maxdim")) { max_dim = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-iters")) { n_iter = (unsigned int)atoi(argv[++i]); continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!flt) flt = argv[i];
        else if (!dists) dists = argv[i];
        else if (!outp) outp = argv[i];
    }
    if (!flt || !dists || !outp) help = true;
    if (help) {
        printf("\nUsage: pr-learn  src_h5_filter_file\n"
               "       src_h5_dist_file dst_h5_output_file\n"
               "       -mu <0.0-1.0, 0.025=default> \n"
               "       -gamma <0.0-10.0, 0.10=default> \n"
               "       -maxdim <64-768, 640=default> \n"
               "       -iters <0-N, 5000000=default> \n\n");
        return 1;
    }
    printf("mu: %g gamma: %g maxdim: %d nIters: %u\n", mu, gamma, max_dim, n_iter);
    // DLCO_MAX_ITERS: stop after that many steps whatever -iters asks for.  workspace/04-prlearn.sh launches
    // `-iters 50000000` and calls a run complete once its log holds the step-5 000 000 line; DLCO_MAX_ITERS=5000000
    // ends a run at exactly that line without touching the script.
    n_iter = max_iters_cap(n_iter);
    try {
        std::vector<float> pr_params;
        {
            h5mini::Reader rf(flt);
            printf("Load PRParams.\n");
            pr_params = rf.read_all<float>("PRParams");
            printf("Load RingParams.\n");
            (void)rf.info("RingParams");                    // read by the reference, not used by it either
        }
        Ctx ctx(0);
        DevBuf<float> d_dists;
        DistFile df = load_dist_file(dists, ctx, d_dists, true);
        const int F = (int)df.f;
        if ((size_t)F * 8 * 3 != pr_params.size()) throw h5mini::Error("PRParams rows must be 8 x the Distance columns");

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
        DevBuf<float> d_w(ctx, F), d_dfavg(ctx, F), d_pd(ctx, df.n);
        std::vector<float> w((size_t)F, 0.0f);
        d_w.upload(w.data(), F);
        d_dfavg.upload(w.data(), F);

        // an existing output keeps its rows: the reference appends (src/pr-learn.cpp:397-407)
        std::vector<float> old_rows;
        if (FILE *f = fopen(outp, "rb")) {
            fclose(f);
            // the reference appends in place (dsinsert); this writer re-creates the file with the old rows in front:
            // anything it cannot carry over is an error, not something to drop silently
            h5mini::Reader ro(outp);
            if (ro.names().size() != 1 || !ro.exists("w"))
                throw h5mini::Error(std::string("existing output holds other datasets than `w` (would be lost): ") + outp);
            if ((int)ro.info("w").row_elems() != F || ro.info("w").dtype != h5mini::F32)
                throw h5mini::Error(std::string("existing output has a `w` of another width: ") + outp);
            old_rows = ro.read_all<float>("w");
        }
        std::unique_ptr<h5mini::Writer> out;
        int hw = -1;
        auto append_w = [&](const float *row) {
            if (!out) {
                out.reset(new h5mini::Writer(outp));
                hw = out->create_chunked("w", h5mini::F32, {(uint64_t)F}, 1, 9, true);
                if (!old_rows.empty()) out->append_rows(hw, old_rows.data(), old_rows.size() / F);
            }
            out->append_rows(hw, row, 1);
            out->flush();
        };

        uint64_t rng_state = 2215;                          // RNG rng( 2215 ), :241
        float obj_best = FLT_MAX;
        int nnz_best = 0;
        double t_train = now_sec();
        uint64_t t = 0;
        while (t <= n_iter) {
            // the reference validates after the update of every step t that is a positive multiple of
            // LogStep (its `step == LogStep` counter, :331)
            uint64_t next_log = (t + log_step - 1) / log_step * log_step;
            if (next_log == 0) next_log = log_step;
            const uint64_t t_end = std::min<uint64_t>(n_iter, next_log);
            check(dlco_pr_rda_run(ctx.h, d_dists.p, F, F, d_pos.p, (int32_t)sp.n_pos_trn, d_neg.p, (int32_t)sp.n_neg_trn, mu,
                                  gamma, t, t_end - t + 1, &rng_state, d_w.p, d_dfavg.p), "dlco_pr_rda_run");
            t = t_end + 1;
            if (t_end != next_log) break;
            ctx.sync();
            const double t_valid = now_sec();
            float loss = 0.f, regul = 0.f;
            check(dlco_pr_validate(ctx.h, d_dists.p, F, F, d_pos.p + sp.n_pos_trn, (int64_t)sp.n_pos_val(),
                                   d_neg.p + sp.n_neg_trn, (int64_t)sp.n_neg_val(), d_w.p, mu, &loss, &regul), "dlco_pr_validate");
            const double t_done = now_sec();
            d_w.download(w.data(), F);
            int nnz = 0;
            for (int i = 0; i < F; i++) nnz += w[i] != 0.0f;
            const float obj = loss + regul;
            if (obj < obj_best) {
                obj_best = obj;
                nnz_best = nnz;
                printf("Best: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f)  NNZ: %i (%i)  Ttime: %.4f Vtime: %.4f\n",
                       (unsigned)t_end, loss, regul, obj, obj_best, nnz, nnz_best, t_valid - t_train, t_done - t_valid);
                const PrDims d = pr_dims(pr_params, w.data(), F);
                float fpr95 = 0.f;
                double auc = 0.0;
                check(dlco_pr_patchdist(ctx.h, d_dists.p, (int64_t)df.n, F, F, d_w.p, d_pd.p), "dlco_pr_patchdist");
                check(dlco_roc_stats(ctx.h, d_pd.p, d_labels.p, (int64_t)df.n, &fpr95, &auc), "dlco_roc_stats");
                if (d.dim <= max_dim) {
                    append_w(w.data());
                    printf("Stat: nPR #%i (#%i) Dim/MaxDim [%i/%i] AUC: %.6f FPR95: %.2f [saved]\n", d.n_pr, d.nz_dim, d.dim,
                           max_dim, auc, fpr95 * 100);
                } else {
                    printf("Stat: nPR #%i (#%i) Dim/MaxDim [%i/%i] AUC: %.6f FPR95: %.2f\n", d.n_pr, d.nz_dim, d.dim, max_dim,
                           auc, fpr95 * 100);
                }
            } else {
                printf("Step: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f)  NNZ: %i (%i)  Ttime: %.4f Vtime: %.4f\n",
                       (unsigned)t_end, loss, regul, obj, obj_best, nnz, nnz_best, t_valid - t_train, t_done - t_valid);
            }
            fflush(stdout);
            t_train = now_sec();
        }
        if (out) out->close();
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npr-learn: %s\n", e.what());
        return 2;
    }
}
