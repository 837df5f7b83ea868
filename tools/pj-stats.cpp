// pj-stats -- FPR95 / AUC of learnt projection models (`W` of pj-learn) on unprojected test
// distance sets.  The reference ships this tool only as a non-building stub (src/pj-stats.cpp: the
// target is commented out in src/CMakeLists.txt:63-64 and the body is a commented copy of
// pr-stats), so its behaviour is DEFINED here from the stub's usage string (:120-124: -maxdim,
// -dst, -prg, -prj), ComputePJStats (src/misc.cpp:266-333) and pr-stats' output format
// (src/pr-stats.cpp:261-266):
//
//   pj-stats [src_h5_filter_file] [-maxdim 80] -dst unproj1.h5 [-dst ...]
//            [-prg pr_model.h5] -prj pj_model1.h5 [[-prg ...] -prj ...]
//
// For every -dst set (a comp-uprjdists file) and every -prj model whose `W` has as many columns as
// the set's `Distance`: Dim = non-zero rows of W, PatchDist = colsum((W Dists^T)^2), ROC.  Models
// with Dim > MaxDim are skipped.  -prg (the pr-learn model a projection was trained on) is only
// echoed; the filter file is accepted and ignored (the unprojected distances already embed it).
#include "common.hpp"

#include <map>

using namespace dlco;

struct ModelStat {
    float fpr95;
    int dim, prj_idx, dst_idx;
};

int main(int argc, char **argv)
{
    int max_dim = 80;
    bool help = false;
    std::vector<std::string> dst_files, prj_files, prg_files;
    std::string pending_prg;
    const char *flt = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-maxdim")) { max_dim = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-dst")) { dst_files.push_back(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prg")) { pending_prg = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prj")) {
                prj_files.push_back(argv[++i]);
                prg_files.push_back(pending_prg);
                pending_prg.clear();
                continue;
            }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!flt) flt = argv[i];
    }
    if (dst_files.empty() || prj_files.empty()) help = true;
    if (help) {
        printf("\nUsage: pj-stats src_h5_filter_file\n"
               "        -maxdim <1-200, 80=default> -dst src_h5_dataset\n"
               "        -prg src_h5_prg_file1 -prj src_h5_prj_file1\n"
               "        -prg src_h5_prg_file2 -prj src_h5_prj_file2\n"
               "        ...\n\n");
        return 1;
    }
    printf("MaxDim: %d nChannels: %d\n", max_dim, 8);
    try {
        Ctx ctx(0);
        std::multimap<double, ModelStat> models;
        for (size_t dst_idx = 0; dst_idx < dst_files.size(); dst_idx++) {
            const std::string &test = dst_files[dst_idx];
            printf("Set: [%s] {full}\n", test.c_str());
            printf("    Load TestLabel: ");
            DevBuf<float> d_dists;
            DistFile df = load_dist_file(test, ctx, d_dists, false);
            const int D = (int)df.f;
            DevBuf<uint8_t> d_labels(ctx, df.n);
            d_labels.upload(df.labels.data(), df.n);
            DevBuf<float> d_W(ctx, (size_t)D * D), d_pd(ctx, df.n);

            std::string model_best;
            double auc_best = 0.0;
            printf("    Validate Model: ");
            Progress bar;
            for (size_t prj_idx = 0; prj_idx < prj_files.size(); prj_idx++) {
                bar.update((double)prj_idx / (double)prj_files.size());
                h5mini::Reader rw(prj_files[prj_idx]);
                const h5mini::DatasetInfo &wi = rw.info("W");
                if (wi.dtype != h5mini::F32 || wi.dims.size() != 2) throw h5mini::Error("W must be a 2-D f32 dataset");
                if ((int)wi.dims[1] != D || (int)wi.dims[0] > D) continue;      // trained on another PR selection
                std::vector<float> W = rw.read_all<float>("W");
                // ComputePJStats keeps the non-zero rows of W (src/misc.cpp:269-277)
                std::vector<float> nz;
                for (uint64_t r = 0; r < wi.dims[0]; r++) {
                    const float *row = &W[(size_t)r * D];
                    bool any = false;
                    for (int c = 0; c < D && !any; c++) any = row[c] != 0.0f;
                    if (any) nz.insert(nz.end(), row, row + D);
                }
                const int dim = (int)(nz.size() / D);
                if (dim > max_dim) continue;
                d_W.upload(nz.data(), nz.size());
                float fpr95 = 0.f;
                double auc = 0.0;
                check(dlco_pj_patchdist(ctx.h, d_dists.p, (int64_t)df.n, D, D, d_W.p, dim, d_pd.p), "dlco_pj_patchdist");
                check(dlco_roc_stats(ctx.h, d_pd.p, d_labels.p, (int64_t)df.n, &fpr95, &auc), "dlco_roc_stats");
                models.insert({auc, ModelStat{fpr95 * 100, dim, (int)prj_idx, (int)dst_idx}});
                if (auc > auc_best) {
                    auc_best = auc;
                    model_best = prj_files[prj_idx];
                }
            }
            bar.update(1.0);
            for (auto &kv : models) {
                const ModelStat &m = kv.second;
                if ((size_t)m.dst_idx != dst_idx) continue;
                const std::string &prg = prg_files[(size_t)m.prj_idx];
                printf("  ModelStat: AUC #%.07g  FPR95: %.2f Dim/MaxDim [%i/%i] nPR: %i (#%i) [%s](#%i)->[%s]%s%s%s\n", kv.first,
                       m.fpr95, m.dim, max_dim, D / 8, D, prj_files[(size_t)m.prj_idx].c_str(), 0, test.c_str(),
                       prg.empty() ? "" : " {", prg.c_str(), prg.empty() ? "" : "}");
            }
            printf("\n  BestModel: [%s] #%i\n", model_best.c_str(), model_best.empty() ? -1 : 0);
            fflush(stdout);
        }
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\npj-stats: %s\n", e.what());
        return 2;
    }
}
