// comp-uprjdists -- drop-in for the reference tool (src/comp-uprjdists.cpp): same argv, datasets
// and stdout lines; the per-pair loop body (:298-335) is dlco_uprjdist_build on blocks of pairs.
//
//   comp-uprjdists src_h5_filter_file src_h5_img_file -prj src_h5_prj_file -id rowid -out dst_h5_file
//
// "Distance" is written with 128-row chunks instead of the reference's {128, 1} column chunks
// (:289-290) -- a storage detail HDF5 readers do not see; h5mini reads both.
#include "common.hpp"

using namespace dlco;

int main(int argc, char **argv)
{
    int widx = 0;
    const int n_bins = 8, b_norm = 1;
    const float sigma = 1.4f;
    bool help = false;
    const char *flt = nullptr, *img = nullptr, *prj = nullptr, *dst = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-prj")) { prj = argv[++i]; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-id")) { widx = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-out")) { dst = argv[++i]; continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!flt) flt = argv[i];
        else if (!img) img = argv[i];
        else help = true;
    }
    if (!flt || !img || !prj || !dst) help = true;
    if (help) {
        printf("\nUsage: comp-uprjdists src_h5_filter_file src_h5_img_file\n"
               "        -prj src_h5_prj_file -id src_h5_prj_matrix_rowid -out dsc_h5_filename\n\n");
        return 1;
    }
    try {
        const int io_threads = host_threads();
        printf("Load PRFilters.\n");
        h5mini::Reader rflt(flt);
        const h5mini::DatasetInfo &fi = rflt.info("PRFilters");
        if (fi.dtype != h5mini::F32 || fi.row_elems() != 4096) throw h5mini::Error("PRFilters must be f32 nF x 64 x 64");
        const int n_filters = (int)fi.dims[0];
        // only the rings the learnt w row selects are needed (SelectPRFilters looks at nothing else): read w first
        // (a missing model file fails here, before any bulk I/O), then those rows of the 671 MB bank
        std::vector<float> w;
        {
            h5mini::Reader rw(prj);
            const h5mini::DatasetInfo &wi = rw.info("w");
            if (wi.dtype != h5mini::F32 || wi.dims.size() != 2 || (int)wi.dims[1] * 8 != n_filters)
                throw h5mini::Error("w must be f32 k x (nFilters / 8)");
            if (widx < 0 || (uint64_t)widx >= wi.dims[0]) throw h5mini::Error("-id is outside the rows of w");
            w.resize(wi.dims[1]);
            rw.read_rows("w", (uint64_t)widx, (uint64_t)widx + 1, w.data());
        }
        std::vector<float> filters((size_t)n_filters * 4096, 0.f);
        {
            Progress bar;
            for (int i = 0; i < n_filters / 8; i++) {
                if (w[i] > 0.0f) rflt.read_rows("PRFilters", (uint64_t)8 * i, (uint64_t)8 * i + 8, &filters[(size_t)8 * i * 4096], io_threads);
                if ((i & 63) == 0) bar.update((double)(8 * i) / (double)n_filters);
            }
            bar.update(1.0);
        }

        printf("ImageSet: [%s]\n", img);
        h5mini::Reader rimg(img);
        printf("Load Indices.\n");
        const h5mini::DatasetInfo &ii = rimg.info("Indices");
        if (ii.dtype != h5mini::I32 || ii.dims.size() != 2 || ii.dims[1] != 4) throw h5mini::Error("Indices must be i32 N x 4");
        std::vector<int32_t> indices = rimg.read_all<int32_t>("Indices", io_threads);
        const int64_t n_pairs = (int64_t)ii.dims[0];
        printf("Load Patches.\n");
        const h5mini::DatasetInfo &pi = rimg.info("Patches");
        if (pi.dtype != h5mini::U8 || pi.dims.size() != 3 || pi.dims[1] != 64 || pi.dims[2] != 64)
            throw h5mini::Error("Patches must be u8 P x 64 x 64");
        const int64_t n_patches = (int64_t)pi.dims[0];
        Ctx ctx(0);
        DevBuf<uint8_t> d_patches(ctx, (size_t)n_patches * 4096);
        {
            Progress bar;
            const int64_t blk = 16384;
            std::vector<uint8_t> buf((size_t)std::min(blk, std::max<int64_t>(n_patches, 1)) * 4096);
            for (int64_t i = 0; i < n_patches; i += blk) {
                const int64_t cnt = std::min(blk, n_patches - i);
                rimg.read_rows("Patches", i, i + cnt, buf.data(), io_threads);
                d_patches.upload(buf.data(), (size_t)cnt * 4096, (size_t)i * 4096);
                bar.update((double)i / (double)n_patches);
            }
            bar.update(1.0);
        }
        for (int64_t i = 0; i < n_pairs; i++)
            if (indices[4 * i] < 0 || indices[4 * i] >= n_patches || indices[4 * i + 2] < 0 || indices[4 * i + 2] >= n_patches)
                throw h5mini::Error("Indices references a patch outside Patches");

        // ---- the learnt w row and SelectPRFilters (src/comp-uprjdists.cpp:222-244) --------------
        printf("Load Learnt Filters: [%s]#%i\n", prj, widx);
        int n_sel = 0;
        std::vector<float> sel = select_pr_filters(filters.data(), n_filters, 4096, w.data(), &n_sel);
        std::vector<float>().swap(filters);
        printf("PRFilters: %i x %i\n", n_sel, 4096);
        printf("Descriptor size: %i\n", n_sel * 8);
        if (n_sel == 0) throw h5mini::Error("the selected w row has no positive weight");
        const int D = n_sel * 8;

        h5mini::Writer out(dst);
        printf("Export Pair Labels: #%lld\n", (long long)n_pairs);
        {
            std::vector<uint8_t> label((size_t)n_pairs);
            for (int64_t i = 0; i < n_pairs; i++) label[i] = indices[4 * i + 1] == indices[4 * i + 3];
            const int h = out.create_chunked("Label", h5mini::U8, {1}, 128, deflate_level(), false);
            Progress bar;
            const int64_t blk = 1 << 16;
            for (int64_t i = 0; i < n_pairs; i += blk) {
                out.append_rows(h, &label[i], (uint64_t)std::min(blk, n_pairs - i), io_threads);
                bar.update((double)(i + blk) / (double)std::max<int64_t>(n_pairs, 1));
            }
            bar.update(1.0);
        }

        dlco_bank *bank = nullptr;
        check(dlco_bank_create(ctx.h, sel.data(), n_sel, &bank), "dlco_bank_create");
        DevBuf<int32_t> d_idx(ctx, (size_t)n_pairs * 4);
        d_idx.upload(indices.data(), (size_t)n_pairs * 4);
        const int hd = out.create_chunked("Distance", h5mini::F32, {(uint64_t)D}, 128, dist_deflate_level(), false);
        announce_storage("comp-uprjdists", dist_deflate_level());

        printf("Start Compute L1 distances.\n");
        const double t0 = now_sec();
        const int64_t block = 32768;
        DevBuf<float> d_out(ctx, (size_t)block * D);
        std::vector<float> h_out((size_t)block * D);
        for (int64_t lo = 0; lo < n_pairs; lo += block) {
            const int64_t hi = std::min(n_pairs, lo + block);
            check(dlco_uprjdist_build(ctx.h, bank, d_patches.p, d_idx.p, lo, hi, n_bins, sigma, b_norm, d_out.p, D),
                  "dlco_uprjdist_build");
            d_out.download(h_out.data(), (size_t)(hi - lo) * D);
            for (size_t i = 0, n = (size_t)(hi - lo) * D; i < n; i++)
                if (!std::isfinite(h_out[i])) {                      // checkRange (:341-345)
                    printf("\nDist contains NaN\n");
                    return -1;
                }
            out.append_rows(hd, h_out.data(), (uint64_t)(hi - lo), io_threads);
            printf("\rStep: %lld / %lld", (long long)hi, (long long)n_pairs);
            fflush(stdout);
        }
        out.close();
        dlco_bank_destroy(ctx.h, bank);
        printf("\nDone.\n\n");
        printf("Total: %.09f sec\n\n", now_sec() - t0);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\ncomp-uprjdists: %s\n", e.what());
        return 2;
    }
}
