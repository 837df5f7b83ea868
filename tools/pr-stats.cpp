// pr-stats -- drop-in for the reference tool (src/pr-stats.cpp): FPR95 / AUC of every row of every
// learnt `w` file on every test distance set; same argv and stdout lines.  ComputePRStats'
// distance pass (w . Dists^T, src/misc.cpp:226) and the ROC (:227-263) run in libdlco_b200.so.
//
//   pr-stats src_h5_filter_file -dst dist1.h5 [-dst ...] -prj model1.h5 [-prj ...]
#include "common.hpp"

#include <map>

using namespace dlco;

struct ModelStat {
    float fpr95;
    int dim, n_pr, nz_dim, prj_idx, dst_idx, widx;
};

int main(int argc, char **argv)
{
    const int max_dim = 640, n_channels = 8;
    bool help = false;
    std::vector<std::string> dst_files, prj_files;
    const char *flt = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-dst")) { dst_files.push_back(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prj")) { prj_files.push_back(argv[++i]); continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!flt) flt = argv[i];
    }
    if (!flt) help = true;
    if (help) {
        printf("\nUsage: pr-stats src_h5_filter_file\n"
               "        -dst src_h5_dst_file1 -dst src_h5_dst_file2 -dst ...\n"
               "        -prj src_h5_prj_file1 -prj src_h5_prj_file2 -prj ...\n\n");
        return 1;
    }
    printf("MaxDim: %d nChannels: %d\n", max_dim, n_channels);
    try {
        std::vector<float> pr_params;
        {
            h5mini::Reader rf(flt);
            printf("Load PRParams.\n");
            pr_params = rf.read_all<float>("PRParams");
            printf("Load RingParams.\n");
            (void)rf.info("RingParams");
        }
        Ctx ctx(0);
        std::multimap<double, ModelStat> models;              // ascending AUC, insertion order among ties
        for (size_t dst_idx = 0; dst_idx < dst_files.size(); dst_idx++) {
            const std::string &test = dst_files[dst_idx];
            printf("Set: [%s] {full}\n", test.c_str());
            printf("    Load TestLabel: ");
            DevBuf<float> d_dists;
            DistFile df = load_dist_file(test, ctx, d_dists, false);
            const int F = (int)df.f;
            if ((size_t)F * 8 * 3 != pr_params.size()) throw h5mini::Error("PRParams rows must be 8 x the Distance columns");
            DevBuf<uint8_t> d_labels(ctx, df.n);
            d_labels.upload(df.labels.data(), df.n);
            DevBuf<float> d_w(ctx, F), d_pd(ctx, df.n);

            std::string model_best;
            int widx_best = -1;
            double auc_best = 0.0;
            printf("    Validate Model: ");
            Progress bar;
            for (size_t prj_idx = 0; prj_idx < prj_files.size(); prj_idx++) {
                h5mini::Reader rw(prj_files[prj_idx]);
                const h5mini::DatasetInfo &wi = rw.info("w");
                if (wi.dtype != h5mini::F32 || wi.dims.size() != 2 || (int)wi.dims[1] != F)
                    throw h5mini::Error("w of " + prj_files[prj_idx] + " does not match the Distance columns");
                std::vector<float> W = rw.read_all<float>("w");
                for (int widx = 0; widx < (int)wi.dims[0]; widx++) {
                    const float *w = &W[(size_t)widx * F];
                    const PrDims d = pr_dims(pr_params, w, F, n_channels);
                    bar.update((double)prj_idx / (double)prj_files.size());
                    if (d.dim > max_dim) continue;             // ComputePRStats' MaxDim abort (:220-222)
                    d_w.upload(w, F);
                    float fpr95 = 0.f;
                    double auc = 0.0;
                    check(dlco_pr_patchdist(ctx.h, d_dists.p, (int64_t)df.n, F, F, d_w.p, d_pd.p), "dlco_pr_patchdist");
                    check(dlco_roc_stats(ctx.h, d_pd.p, d_labels.p, (int64_t)df.n, &fpr95, &auc), "dlco_roc_stats");
                    models.insert({auc, ModelStat{fpr95 * 100, d.dim, d.n_pr, d.nz_dim, (int)prj_idx, (int)dst_idx, widx}});
                    if (auc > auc_best) {
                        auc_best = auc;
                        widx_best = widx;
                        model_best = prj_files[prj_idx];
                    }
                }
            }
            bar.update(1.0);
            for (auto &kv : models) {
                const ModelStat &m = kv.second;
                if (dst_files[(size_t)m.dst_idx] != test) continue;
                printf("  ModelStat: AUC #%.07g  FPR95: %.2f Dim/MaxDim [%i/%i] nPR: %i (#%i) [%s](#%i)->[%s]\n", kv.first,
                       m.fpr95, m.dim, max_dim, m.n_pr, m.nz_dim, prj_files[(size_t)m.prj_idx].c_str(), m.widx,
                       dst_files[(size_t)m.dst_idx].c_str());
            }
            printf("\n  BestModel: [%s] #%i\n", model_best.c_str(), widx_best);
            fflush(stdout);
        }
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npr-stats: %s\n", e.what());
        return 2;
    }
}
