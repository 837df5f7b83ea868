// gen-poolregion -- drop-in for the reference tool (src/gen-poolregion.cpp): same argv, stdout lines
// and datasets ("RingParams" f64 n x 3, "PRParams" f32 8n x 3, "PRFilters" f32 8n x P x P, chunks of
// 128 rows, deflate).  The ring grid is dlco_ring_params (host loops as written in the reference,
// :163-213), the filters come from the device (dlco_gen_pr_filters, one CTA per filter, fp64 math,
// bit-identical to the reference's libm evaluation) in blocks that are deflated and written while
// the next block is generated.
//
//   gen-poolregion dst_h5_file [-patchsize 64] [-nr 32] [-nphi 5] [-nsigma 32]
#include "common.hpp"

using namespace dlco;

int main(int argc, char **argv)
{
    int nr = 32, nphi = 5, nsigma = 32, patch = 64;
    const int sChunk = 128;
    bool help = false;
    const char *dst = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-patchsize")) { patch = (uint8_t)atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-nr")) { nr = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-nphi")) { nphi = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-nsigma")) { nsigma = atoi(argv[++i]); continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!dst) dst = argv[i];
    }
    if (!dst) help = true;
    if (help) {
        printf("\nUsage: gen-poolregion dst_h5_file\n"
               "       -patchsize <0-255, 64=default> \n"
               "       -nr <0-255, 32=default> \n"
               "       -nphi <0-255, 5=default> \n"
               "       -nsigma <0-255, 32=default> \n\n");
        return 1;
    }
    try {
        if (patch != DLCO_PATCH) throw h5mini::Error("only -patchsize 64 is supported (the trainer's kernels are built for 64 x 64 patches)");
        const int io_threads = host_threads();
        int n_rings = 0;
        check(dlco_ring_params(patch, nr, nphi, nsigma, nullptr, 0, &n_rings), "dlco_ring_params");
        // the reference sizes its matrices by nr*nphi*nsigma and fills them from the double loops; they agree for
        // the defaults (5120).  Where the accumulated loop variables yield fewer rows the rest stays zero there:
        // reproduce the row count, refuse the overflow case (the reference writes out of bounds)
        const int n_params = nr * nphi * nsigma;
        if (n_rings > n_params) throw h5mini::Error("the ring loops yield more rows than nr*nphi*nsigma");
        std::vector<double> rings((size_t)n_params * 3, 0.0);
        check(dlco_ring_params(patch, nr, nphi, nsigma, rings.data(), n_params, &n_rings), "dlco_ring_params");

        Ctx ctx(0);                                           // before the output file exists: no GPU, no partial file
        h5mini::Writer out(dst);
        printf("Export RingParams: #%d\n", n_params);
        {
            const int h = out.create_chunked("RingParams", h5mini::F64, {3}, sChunk, deflate_level(), false);
            Progress bar;
            out.append_rows(h, rings.data(), (uint64_t)n_params, io_threads);
            bar.update(1.0);
        }
        const int nPR = n_params * 8;
        printf("Export Projection: #%d\n", nPR);
        const int hp = out.create_chunked("PRParams", h5mini::F32, {3}, sChunk, deflate_level(), false);
        const int hf = out.create_chunked("PRFilters", h5mini::F32, {(uint64_t)patch, (uint64_t)patch}, sChunk, deflate_level(), false);
        const int block_rings = 256;                          // 2048 filters = 32 MB per block
        DevBuf<float> d_f(ctx, (size_t)block_rings * 8 * 4096);
        std::vector<float> h_f((size_t)block_rings * 8 * 4096), h_p((size_t)n_params * 8 * 3, 0.f);
        Progress bar;
        int fixed_total = 0;
        for (int r0 = 0; r0 < n_params; r0 += block_rings) {
            const int cnt = std::min(block_rings, n_params - r0);
            int fixed = 0;
            // rows beyond n_rings (none for the defaults) are all-zero rings: sigma = 0 is not a filter
            const int live = std::max(0, std::min(cnt, n_rings - r0));
            std::fill(h_f.begin(), h_f.end(), 0.f);
            if (live > 0) {
                check(dlco_gen_pr_filters(ctx.h, &rings[(size_t)r0 * 3], live, d_f.p, &h_p[(size_t)r0 * 8 * 3], &fixed),
                      "dlco_gen_pr_filters");
                d_f.download(h_f.data(), (size_t)live * 8 * 4096);
            }
            fixed_total += fixed;
            out.append_rows(hf, h_f.data(), (uint64_t)cnt * 8, io_threads);
            bar.update((double)r0 / (double)n_params);
        }
        out.append_rows(hp, h_p.data(), (uint64_t)nPR, io_threads);
        bar.update(1.0);
        out.close();
        if (fixed_total) fprintf(stderr, "gen-poolregion: %d filters regenerated with the host libm (rounding ties)\n", fixed_total);
        printf("Done.\n\n");
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\ngen-poolregion: %s\n", e.what());
        return 2;
    }
}
