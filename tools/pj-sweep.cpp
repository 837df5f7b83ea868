// pj-sweep -- the (mu, gamma) loop of workspace/08-pjlearn.sh on one data set as ONE process per GPU.
//
// The reference runs `pj-learn -iters N -mu m -gamma g <ds>-unproj.h5 <ds>-m-g-pj.h5 | tee
// logging/<ds>-m-g-pj.log` once per grid point (9 x 45 runs).  Every run shuffles with the default
// theRNG() and samples with RNG(2215) (src/pj-learn.cpp:214-227), so all grid points of one data
// set draw identical batches: this tool advances them in lock step through
// dlco_pj_rda_step_grid (one gather, one grouped SYRK, one batched eigen-decomposition per
// iteration) and writes, per grid point, exactly the files the loop would have produced -- the
// model `<outdir>/<ds>-<mu>-<gamma>-pj.h5` (W, A) and the log `<outdir>/logging/<ds>-<mu>-<gamma>-pj.log`
// with pj-learn's stdout lines -- so 09-pjstats.sh / 10-pj-top.sh read them unchanged.
//
//   pj-sweep src_h5_dist_file out_dir -mu m1,m2,.. -gamma g1,g2,.. [-iters n] [-ds name]
//            [-gpu i] [-shard r/n] [-prox eig|sign]
// -shard r/n: this process owns grid points c with c % n == r (one process per GPU of the box).
// -prox eig (default): the reference's prox, an eigen-decomposition per iteration (src/pj-learn.cpp:434-490).
// -prox sign: the projection by the tensor-core matrix-sign iteration (dlco_pj_rda_step_grid_sign); W is
//             rebuilt from the projected A (dlco_pj_factor_grid) at the validation points only.
#include "common.hpp"

#include <cfloat>
#include <sys/stat.h>

using namespace dlco;

static std::vector<std::string> split_list(const char *s)
{
    std::vector<std::string> out;
    std::string cur;
    for (const char *p = s;; p++) {
        if (*p == ',' || *p == 0) {
            if (!cur.empty()) out.push_back(cur);
            cur.clear();
            if (!*p) break;
        } else cur.push_back(*p);
    }
    return out;
}

struct Cfg {
    std::string mu_s, gamma_s;
    float mu, gamma;
    FILE *log = nullptr;
    float obj_best = FLT_MAX, fpr95_best = 1.0f;
    double auc_best = 0.0;
    int rows_best = 0, rank_save = -1;
    std::vector<float> W_save, A_save;
};

int main(int argc, char **argv)
{
    unsigned int n_iter = 50000;
    const unsigned int log_step = 100, sz_batch = 200;
    int gpu = 0, shard_r = 0, shard_n = 1;
    bool prox_sign = false;
    bool help = false;
    const char *dists = nullptr, *outdir = nullptr, *mus = nullptr, *gammas = nullptr, *ds_name = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-mu")) { mus = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gamma")) { gammas = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-iters")) { n_iter = (unsigned int)atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-ds")) { ds_name = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gpu")) { gpu = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prox")) {
                const char *v = argv[++i];
                if (!strcmp(v, "sign")) prox_sign = true;
                else if (!strcmp(v, "eig")) prox_sign = false;
                else { printf("ERROR: Invalid -prox value (eig | sign).\n"); return 1; }
                continue;
            }
            if (i + 1 < argc && !strcmp(argv[i], "-shard")) {
                if (sscanf(argv[++i], "%d/%d", &shard_r, &shard_n) != 2 || shard_n < 1 || shard_r < 0 || shard_r >= shard_n) {
                    printf("ERROR: -shard wants r/n with 0 <= r < n.\n");
                    help = true;
                }
                continue;
            }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!dists) dists = argv[i];
        else if (!outdir) outdir = argv[i];
    }
    if (!dists || !outdir || !mus || !gammas) help = true;
    if (help) {
        printf("\nUsage: pj-sweep src_h5_dist_file out_dir -mu m1,m2,... -gamma g1,g2,...\n"
               "       -iters <0-N, 50000=default> -ds <name, default: file stem without -unproj>\n"
               "       -gpu <device, 0=default> -shard <r/n, 0/1=default>\n\n");
        return 1;
    }
    std::string ds;
    if (ds_name) ds = ds_name;
    else {
        ds = dists;
        size_t sl = ds.find_last_of('/');
        if (sl != std::string::npos) ds = ds.substr(sl + 1);
        size_t ext = ds.rfind(".h5");
        if (ext != std::string::npos) ds = ds.substr(0, ext);
        size_t un = ds.rfind("-unproj");
        if (un != std::string::npos) ds = ds.substr(0, un);
    }
    // grid order of the script: mu outer, gamma inner
    std::vector<Cfg> cfgs;
    {
        int c = 0;
        for (const std::string &m : split_list(mus))
            for (const std::string &g : split_list(gammas)) {
                if (c++ % shard_n != shard_r) continue;
                Cfg k;
                k.mu_s = m;
                k.gamma_s = g;
                k.mu = (float)atof(m.c_str());
                k.gamma = (float)atof(g.c_str());
                cfgs.push_back(std::move(k));
            }
    }
    const int n_cfg = (int)cfgs.size();
    if (n_cfg == 0) { printf("pj-sweep: no grid point for shard %d/%d\n", shard_r, shard_n); return 0; }
    const std::string od = outdir, ld = od + "/logging";
    mkdir(od.c_str(), 0777);
    mkdir(ld.c_str(), 0777);
    printf("pj-sweep: data set [%s], %d grid points on GPU %d (shard %d/%d), nIters: %u\n", ds.c_str(), n_cfg, gpu, shard_r,
           shard_n, n_iter);
    try {
        Ctx ctx(gpu);
        DevBuf<float> d_dists;
        DistFile df = load_dist_file(dists, ctx, d_dists, true);
        const int D = (int)df.f;
        Split sp = split_train_val(df.labels);
        if (!sp.n_pos_trn || !sp.n_neg_trn) throw h5mini::Error("need positive and negative training pairs");
        char gname[256];
        int ma = 0, mi = 0;
        check(dlco_device_name(ctx.h, gname, sizeof(gname), &ma, &mi), "dlco_device_name");
        for (Cfg &k : cfgs) {                                // the header pj-learn prints (src/pj-learn.cpp:166-272)
            const std::string lp = ld + "/" + ds + "-" + k.mu_s + "-" + k.gamma_s + "-pj.log";
            k.log = fopen(lp.c_str(), "w");
            if (!k.log) throw h5mini::Error("cannot write " + lp);
            fprintf(k.log, "mu: %g gamma: %g nIters: %u\n", k.mu, k.gamma, n_iter);
            fprintf(k.log, "Load Labels: %llu\nLoad Distances: %llu x %llu\n", (unsigned long long)df.n,
                    (unsigned long long)df.n, (unsigned long long)df.f);
            fprintf(k.log, "Positive samples #%zu\nNegative samples #%zu\n", sp.idx_pos.size(), sp.idx_neg.size());
            fprintf(k.log, "Positive train #%zu\nNegative train #%zu\nPositive valid #%zu\nNegative valid #%zu\n", sp.n_pos_trn,
                    sp.n_neg_trn, sp.n_pos_val(), sp.n_neg_val());
            fprintf(k.log, "\nFound GPU: %s\nCompute Capability: %d.%d\n\n", gname, ma, mi);
        }

        DevBuf<int32_t> d_pos(ctx, sp.idx_pos.size()), d_neg(ctx, sp.idx_neg.size());
        d_pos.upload(sp.idx_pos.data(), sp.idx_pos.size());
        d_neg.upload(sp.idx_neg.data(), sp.idx_neg.size());
        DevBuf<uint8_t> d_labels(ctx, df.n);
        d_labels.upload(df.labels.data(), df.n);
        const size_t DD = (size_t)D * D;
        DevBuf<float> d_W(ctx, n_cfg * DD), d_A(ctx, n_cfg * DD), d_dfavg(ctx, n_cfg * DD);
        DevBuf<float> d_work(ctx, n_cfg * (2 * DD + D)), d_mu(ctx, n_cfg), d_gamma(ctx, n_cfg), d_pd(ctx, df.n);
        DevBuf<int32_t> d_rank(ctx, n_cfg);
        DevBuf<uint64_t> d_rng(ctx, 1);
        {
            std::vector<float> z(n_cfg * DD, 0.0f), m(n_cfg), g(n_cfg);
            d_W.upload(z.data(), z.size());
            d_A.upload(z.data(), z.size());
            d_dfavg.upload(z.data(), z.size());
            for (int c = 0; c < n_cfg; c++) { m[c] = cfgs[c].mu; g[c] = cfgs[c].gamma; }
            d_mu.upload(m.data(), n_cfg);
            d_gamma.upload(g.data(), n_cfg);
            std::vector<int32_t> r0(n_cfg, 0);               // W = zeros(D, D): all margins 0 at t = 0
            d_rank.upload(r0.data(), n_cfg);
            const uint64_t seed = 2215;                      // RNG rng( 2215 ), :225
            d_rng.upload(&seed, 1);
        }
        std::vector<int32_t> ranks(n_cfg);
        unsigned int step = 0;
        double t_train = now_sec();
        const double t_all = now_sec();
        double t_factor = 0.0, t_validate = 0.0, t_stats = 0.0;      // where the validation points spend their time
        n_iter = max_iters_cap(n_iter);                    // DLCO_MAX_ITERS, see tools/common.hpp
        for (unsigned int t = 0; t <= n_iter; t++) {
            if (prox_sign)
                check(dlco_pj_rda_step_grid_sign(ctx.h, d_dists.p, D, D, d_pos.p, (int32_t)sp.n_pos_trn, d_neg.p,
                                                 (int32_t)sp.n_neg_trn, (int)sz_batch, d_rng.p, n_cfg, d_mu.p, d_gamma.p, t,
                                                 d_A.p, d_dfavg.p, d_rank.p, d_work.p),
                      "dlco_pj_rda_step_grid_sign");
            else
                check(dlco_pj_rda_step_grid(ctx.h, d_dists.p, D, D, d_pos.p, (int32_t)sp.n_pos_trn, d_neg.p, (int32_t)sp.n_neg_trn,
                                            (int)sz_batch, d_rng.p, n_cfg, d_mu.p, d_gamma.p, t, d_W.p, d_A.p, d_dfavg.p, d_rank.p,
                                            d_work.p),
                      "dlco_pj_rda_step_grid");
            if (step == log_step) {
                ctx.sync();                                    // the iterations queued so far: training time, not validation
                const double t_f = now_sec();
                if (prox_sign)      // W of the projected A for validation, statistics and the saved model (synchronises)
                    check(dlco_pj_factor_grid(ctx.h, D, n_cfg, d_A.p, d_rank.p, d_W.p, (int64_t)DD, d_work.p), "dlco_pj_factor_grid");
                else
                    check(dlco_pj_grid_check(ctx.h, n_cfg), "dlco_pj_grid_check");      // synchronises
                d_rank.download(ranks.data(), n_cfg);
                t_factor += now_sec() - t_f;
                const double ttime = now_sec() - t_train;
                for (int c = 0; c < n_cfg; c++) {
                    Cfg &k = cfgs[c];
                    const int rank = ranks[c];
                    const float *Wc = d_W.p + c * DD, *Ac = d_A.p + c * DD;
                    const double t_valid = now_sec();
                    float loss = 0.f, regul = 0.f;
                    check(dlco_pj_validate(ctx.h, d_dists.p, D, D, d_pos.p + sp.n_pos_trn, (int64_t)sp.n_pos_val(),
                                           d_neg.p + sp.n_neg_trn, (int64_t)sp.n_neg_val(), Wc, rank, Ac, k.mu, &loss, &regul),
                          "dlco_pj_validate");
                    const double vtime = now_sec() - t_valid;
                    t_validate += vtime;
                    const float obj = loss + regul;
                    const int w_rows = rank > 0 ? rank : D;
                    if (obj < k.obj_best) {
                        k.obj_best = obj;
                        k.rows_best = w_rows;
                        fprintf(k.log, "Best: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t,
                                loss, regul, obj, k.obj_best, w_rows, k.rows_best, ttime, vtime);
                        float fpr95 = 0.f;
                        double auc = 0.0;
                        const double t_s = now_sec();
                        check(dlco_pj_patchdist(ctx.h, d_dists.p, (int64_t)df.n, D, D, Wc, rank, d_pd.p), "dlco_pj_patchdist");
                        check(dlco_roc_stats(ctx.h, d_pd.p, d_labels.p, (int64_t)df.n, &fpr95, &auc), "dlco_roc_stats");
                        const bool keep = k.auc_best <= auc && k.fpr95_best >= fpr95;       // :560-566
                        if (keep) {
                            k.auc_best = auc;
                            k.fpr95_best = fpr95;
                            k.rank_save = w_rows;
                            k.W_save.resize((size_t)w_rows * D);
                            k.A_save.resize(DD);
                            d_W.download(k.W_save.data(), k.W_save.size(), c * DD);
                            d_A.download(k.A_save.data(), DD, c * DD);
                        }
                        t_stats += now_sec() - t_s;
                        fprintf(k.log, "Stat: Dim [%i] AUC: %.6f (%.6f) FPR95: %.2f (%.2f)%s\n", rank, auc, k.auc_best, fpr95 * 100,
                                k.fpr95_best * 100, keep ? " [saved]" : "");
                    } else {
                        fprintf(k.log, "Step: %u  Loss: %.6f Regul: %.6f Obj: %.6f (%.6f) Rank: %i (%i) Ttime: %.4f Vtime: %.4f\n", t,
                                loss, regul, obj, k.obj_best, w_rows, k.rows_best, ttime, vtime);
                    }
                    fflush(k.log);
                }
                step = 0;
                t_train = now_sec();
            }
            step++;
        }
        ctx.sync();
        const double secs = now_sec() - t_all;
        for (Cfg &k : cfgs) {                                // W_Save / A_Save (src/pj-learn.cpp:592-594)
            if (k.rank_save < 0) {
                k.rank_save = D;
                k.W_save.assign(DD, 0.0f);
                k.A_save.assign(DD, 0.0f);
            }
            const std::string op = od + "/" + ds + "-" + k.mu_s + "-" + k.gamma_s + "-pj.h5";
            h5mini::Writer out(op);
            out.write_contiguous("W", h5mini::F32, {(uint64_t)k.rank_save, (uint64_t)D}, k.W_save.data());
            out.write_contiguous("A", h5mini::F32, {(uint64_t)D, (uint64_t)D}, k.A_save.data());
            out.close();
            fclose(k.log);
            k.log = nullptr;
        }
        printf("pj-sweep: %d grid points x %u iterations in %.2f s (%.1f config-iterations/s)\n", n_cfg, n_iter + 1, secs,
               (double)n_cfg * (n_iter + 1) / secs);
        printf("pj-sweep: validation points took %.2f s (factor / sync %.2f, validate %.2f, statistics + model copies %.2f)\n",
               t_factor + t_validate + t_stats, t_factor, t_validate, t_stats);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npj-sweep: %s\n", e.what());
        return 2;
    }
}
