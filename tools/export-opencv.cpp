// export-opencv -- drop-in for the reference tool (src/export-opencv.cpp): writes the learnt pooling regions
// (SelectPRFilters of a `w` row) and the projection `W` as the sparse C tables OpenCV's xfeatures2d VGG
// descriptor includes (workspace/opencv/vgg_generated_*.i).  Host only: no arithmetic, the float bit
// patterns are printed as they are stored.
//
//   export-opencv -flt filters.h5 -prg pr.h5 -id row -prj pj.h5 out_header.i
#include "common.hpp"

using namespace dlco;

// run-length index ("start,count" of every run of non-zeros, row-major) -- src/export-opencv.cpp:232-270 and
// :314-352.  The reference leaves the inner loop with `break` once the last non-zero is out; when that happens
// before the last row, the following rows re-enter with the run still open and emit it again -- kept as is.
static void write_idx(FILE *out, const float *m, int rows, int cols, int total)
{
    int idx = 0, start = -1, count = 0;
    for (int r = 0; r < rows; r++)
        for (int c = 0; c < cols; c++) {
            const bool nz = m[(size_t)r * cols + c] != 0.0f;
            if (nz) {
                total--;
                count++;
                if (start == -1) start = r * cols + c;
            }
            if (total != 0 && nz) continue;
            if (count == 0) continue;
            idx++;
            if (idx == 1) fprintf(out, "{\n ");
            fprintf(out, "0x%x,0x%X", start, count);
            if (total == 0) {
                fprintf(out, "\n};\n");
                break;
            }
            fprintf(out, ",");
            if (idx % 8 == 0) fprintf(out, "\n ");
            start = -1;
            count = 0;
        }
}

// the non-zero elements as 32-bit patterns, 8 per line -- src/export-opencv.cpp:277-301 and :359-383
static void write_vals(FILE *out, const float *m, int rows, int cols, int total)
{
    int count = 1;
    for (int r = 0; r < rows; r++)
        for (int c = 0; c < cols; c++) {
            if (r + c == 0) fprintf(out, "{\n ");
            const float v = m[(size_t)r * cols + c];
            if (v == 0.0f) continue;
            uint32_t bits;
            memcpy(&bits, &v, 4);
            fprintf(out, "0x%08x", bits);
            if (count == total) {
                fprintf(out, "\n};\n");
                break;
            }
            fprintf(out, ",");
            if (count % 8 == 0) fprintf(out, "\n ");
            count++;
        }
}

static int count_nonzero(const std::vector<float> &m)
{
    int n = 0;
    for (float v : m) n += v != 0.0f;
    return n;
}

int main(int argc, char **argv)
{
    int widx = -1;
    bool help = false;
    const char *flt = nullptr, *prg = nullptr, *prj = nullptr, *dst = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-flt")) { flt = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prg")) { prg = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-id")) { widx = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prj")) { prj = argv[++i]; continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!dst) dst = argv[i];
        else help = true;
    }
    if (widx == -1 || !flt || !prg || !prj || !dst) help = true;
    if (help) {
        printf("\nUsage: export-opencv -flt src_h5_filter_file\n"
               "       -prg src_h5_prg_file -id src_h5_prg_matrix_rowid\n"
               "       -prj src_h5_prj_file out_cc_headerfile\n\n");
        return 1;
    }
    try {
        const int io_threads = host_threads();
        h5mini::Reader rflt(flt);
        const h5mini::DatasetInfo &fi = rflt.info("PRFilters");
        if (fi.dtype != h5mini::F32 || fi.dims.size() != 3) throw h5mini::Error("PRFilters must be f32 nF x P x P");
        const int n_filters = (int)fi.dims[0], fsize = (int)fi.row_elems();
        printf("Load PRFilters:\n");
        std::vector<float> w;
        {
            h5mini::Reader rw(prg);
            const h5mini::DatasetInfo &wi = rw.info("w");
            if (wi.dtype != h5mini::F32 || wi.dims.size() != 2 || (int)wi.dims[1] * 8 != n_filters)
                throw h5mini::Error("w must be f32 k x (nFilters / 8)");
            if ((uint64_t)widx >= wi.dims[0]) throw h5mini::Error("-id is outside the rows of w");
            w.resize(wi.dims[1]);
            rw.read_rows("w", (uint64_t)widx, (uint64_t)widx + 1, w.data());
        }
        // only the rings with w > 0 can be selected: read those rows, leave the rest zero (the reference reads all
        // 671 MB; SelectPRFilters looks at nothing else)
        std::vector<float> filters((size_t)n_filters * fsize, 0.f);
        {
            Progress bar;
            for (int i = 0; i < n_filters / 8; i++) {
                if (w[i] > 0.0f) rflt.read_rows("PRFilters", (uint64_t)8 * i, (uint64_t)8 * i + 8, &filters[(size_t)8 * i * fsize], io_threads);
                bar.update((double)(8 * i) / (double)n_filters);
            }
            bar.update(1.0);
        }
        printf("Load Learnt Filters: [%s]#%i\n", prg, widx);
        printf("Load Learnt Projections: [%s]\n", prj);
        h5mini::Reader rp(prj);
        const h5mini::DatasetInfo &Wi = rp.info("W");
        if (Wi.dtype != h5mini::F32 || Wi.dims.size() != 2) throw h5mini::Error("W must be a 2-D f32 dataset");
        const int Wrows = (int)Wi.dims[0], Wcols = (int)Wi.dims[1];
        std::vector<float> W = rp.read_all<float>("W", io_threads);

        int n_sel = 0;
        std::vector<float> sel = select_pr_filters(filters.data(), n_filters, fsize, w.data(), &n_sel);
        printf("PRFilters: %i x %i [%i]\n", n_sel, fsize, n_sel * 8);
        printf("PJFilters: %i x [%i]\n", Wrows, Wcols);
        if (Wcols != n_sel * 8) {
            printf("ERROR: PJFilters [%i] not agree PRFilters [%i].\n", n_sel * 8, Wcols);
            return 0;
        }
        FILE *out = fopen(dst, "w");
        if (!out) throw h5mini::Error(std::string("cannot create ") + dst);
        fprintf(out, "// generated VGG pooling region filters & projection parameters\n\n");
        fprintf(out, "// PR: [%s]#%i\n// PJ: [%s]\n\n\n", prg, widx, prj);
        fprintf(out, "// PR orig rows\nstatic const int PRrows = %i;\n\n", n_sel);
        fprintf(out, "// PR orig cols\nstatic const int PRcols = %i;\n\n", fsize);
        fprintf(out, "// PR indexes & len\nstatic const unsigned int PRidx[] =\n");
        write_idx(out, sel.data(), n_sel, fsize, count_nonzero(sel));
        fprintf(out, "\n// PR matrix\nstatic const unsigned int PR[] =\n");
        write_vals(out, sel.data(), n_sel, fsize, count_nonzero(sel));
        fprintf(out, "\n\n// PJ orig rows\nstatic const int PJrows = %i;\n\n", Wrows);
        fprintf(out, "// PJ orig cols\nstatic const int PJcols = %i;\n\n", Wcols);
        fprintf(out, "// PJ indexes & len\nstatic const unsigned int PJidx[] =\n");
        write_idx(out, W.data(), Wrows, Wcols, count_nonzero(W));
        fprintf(out, "\n// PJ sparse elements\nstatic const unsigned int PJ[] =\n");
        write_vals(out, W.data(), Wrows, Wcols, count_nonzero(W));
        fclose(out);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\nexport-opencv: %s\n", e.what());
        return 2;
    }
}
