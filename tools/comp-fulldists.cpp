// comp-fulldists -- drop-in for the reference tool of the same name (src/comp-fulldists.cpp): same
// argv, same HDF5 datasets in and out, same stdout lines; the per-pair loop body (:294-356) is one
// call into libdlco_b200.so per block of pairs.
//
//   comp-fulldists src_h5_filter_file src_h5_patches_file dst_h5_dist_file
//                  [-anglebins 8] [-sigma 1.4] [-norm 1] [-gpus N]
//
// -gpus N (extension, default 1): blocks of pairs are dealt round-robin to N GPUs, one host thread
// and one dlco_ctx per GPU; rows are written in pair order.
#include "common.hpp"

#include <atomic>
#include <memory>
#include <condition_variable>
#include <mutex>

#include <cuda_runtime_api.h>

using namespace dlco;

static void usage()
{
    printf("\nUsage: comp-fulldists src_h5_filter_file src_h5_patches_file dst_h5_dist_file\n"
           "       -anglebins <0-255, 8=default> \n"
           "       -sigma <0.0-10.0, 1.4=default> \n"
           "       -norm <0,1 1=default> \n\n");
}

struct Block {
    std::mutex m;
    std::condition_variable cv;
    bool ready = false, free_ = true;
    float *host = nullptr;      // pinned, block_pairs x rings
    int64_t lo = 0, hi = 0;
    std::string error;
};

int main(int argc, char **argv)
{
    int n_bins = 8, b_norm = 1, n_gpus = 1;
    float sigma = 1.4f;
    bool help = false;
    const char *flt = nullptr, *img = nullptr, *dst = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-anglebins")) { n_bins = atoi(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-sigma")) { sigma = (float)atof(argv[++i]); continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-norm")) { b_norm = atoi(argv[++i]) != 0; continue; }
            if (i + 1 < argc && !strcmp(argv[i], "-gpus")) { n_gpus = std::max(1, atoi(argv[++i])); continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!flt) flt = argv[i];
        else if (!img) img = argv[i];
        else if (!dst) dst = argv[i];
    }
    if (!flt || !img || !dst) help = true;
    if (help) { usage(); return 1; }
    printf("Bins: #%d Sigma: %g bNorm: %d\n", n_bins, sigma, b_norm);

    try {
        const int io_threads = host_threads();
        // ---- inputs (src/comp-fulldists.cpp:150-222) ------------------------------------------
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
        std::vector<uint8_t> patches((size_t)n_patches * 4096);
        {
            Progress bar;
            const int64_t blk = 16384;
            for (int64_t i = 0; i < n_patches; i += blk) {
                rimg.read_rows("Patches", i, std::min(n_patches, i + blk), &patches[(size_t)i * 4096], io_threads);
                bar.update((double)i / (double)n_patches);
            }
            bar.update(1.0);
        }

        printf("Load PRFilters.\n");
        h5mini::Reader rflt(flt);
        const h5mini::DatasetInfo &fi = rflt.info("PRFilters");
        if (fi.dtype != h5mini::F32 || fi.row_elems() != 4096) throw h5mini::Error("PRFilters must be f32 nF x 64 x 64");
        const int n_filters = (int)fi.dims[0];
        if (n_filters % 8) throw h5mini::Error("PRFilters: the number of filters must be a multiple of 8");
        std::vector<float> filters((size_t)n_filters * 4096);
        {
            Progress bar;
            const int blk = 2048;
            for (int i = 0; i < n_filters; i += blk) {
                rflt.read_rows("PRFilters", i, std::min(n_filters, i + blk), &filters[(size_t)i * 4096], io_threads);
                bar.update((double)i / (double)n_filters);
            }
            bar.update(1.0);
        }
        const int rings = n_filters / 8;

        // ---- labels (src/comp-fulldists.cpp:225-264) ------------------------------------------
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

        // ---- one context + bank per GPU ---------------------------------------------------------
        std::vector<std::unique_ptr<Ctx>> ctxs;
        std::vector<dlco_bank *> banks((size_t)n_gpus, nullptr);
        for (int g = 0; g < n_gpus; g++) ctxs.emplace_back(new Ctx(g));
        ctxs[0]->print_gpu();
        for (int g = 0; g < n_gpus; g++) check(dlco_bank_create(ctxs[g]->h, filters.data(), n_filters, &banks[g]), "dlco_bank_create");
        std::vector<float>().swap(filters);

        const int hd = out.create_chunked("Distance", h5mini::F32, {(uint64_t)rings}, 128, dist_deflate_level(), false);
        announce_storage("comp-fulldists", dist_deflate_level());
        printf("Start Compute L2 distances.\n");
        const double t0 = now_sec();

        // ---- the hot loop: blocks of pairs, two in flight per GPU, written in order -----------------
        // Each GPU's worker queues block k with the non-draining entry point and only then waits for block k-1, so the
        // device never idles between blocks while the writer thread deflates and writes finished blocks in order.
        const int64_t block_pairs = 8192;
        const int64_t n_blocks = (n_pairs + block_pairs - 1) / block_pairs;
        const int per_gpu = 3;                                   // computing, queued behind it, being written
        const int slots = per_gpu * n_gpus;
        std::vector<Block> blk((size_t)slots);
        for (auto &b : blk)
            if (cudaMallocHost((void **)&b.host, (size_t)block_pairs * rings * sizeof(float)) != cudaSuccess)
                throw AbiError("cudaMallocHost failed for the output staging buffers");
        std::atomic<bool> failed(false);
        auto fail_all = [&](const std::string &why) {          // wake everybody; every wait predicate looks at `failed`
            failed = true;
            for (auto &x : blk) {
                std::lock_guard<std::mutex> lk(x.m);
                if (x.error.empty()) x.error = why;
                x.cv.notify_all();
            }
        };
        std::vector<std::thread> workers;
        for (int g = 0; g < n_gpus; g++)
            workers.emplace_back([&, g]() {
                int64_t prev_ticket = -1;
                Block *prev = nullptr;
                auto finish_prev = [&]() -> bool {
                    if (!prev) return true;
                    const int rc = dlco_fulldist_host_wait(ctxs[g]->h, prev_ticket);
                    if (rc != DLCO_OK) { fail_all(dlco_last_error()); return false; }
                    std::lock_guard<std::mutex> lk(prev->m);
                    prev->ready = true;
                    prev->cv.notify_all();
                    prev = nullptr;
                    return true;
                };
                for (int64_t b = g, k = 0; b < n_blocks && !failed; b += n_gpus, k++) {
                    Block &s = blk[(size_t)(g * per_gpu + (k % per_gpu))];
                    {
                        std::unique_lock<std::mutex> lk(s.m);
                        s.cv.wait(lk, [&] { return s.free_ || failed.load(); });
                        if (failed) break;
                        s.free_ = false;
                    }
                    s.lo = b * block_pairs;
                    s.hi = std::min(n_pairs, s.lo + block_pairs);
                    int64_t ticket = -1;
                    const int rc = dlco_fulldist_build_host_async(ctxs[g]->h, banks[g], patches.data(), n_patches, indices.data(),
                                                                  s.lo, s.hi, n_bins, sigma, b_norm, s.host, rings, &ticket);
                    if (rc != DLCO_OK) { fail_all(dlco_last_error()); break; }
                    if (!finish_prev()) break;
                    prev = &s;
                    prev_ticket = ticket;
                }
                if (!failed) finish_prev();
                dlco_sync(ctxs[g]->h);
            });
        std::string error;
        for (int64_t b = 0; b < n_blocks && error.empty(); b++) {
            const int g = (int)(b % n_gpus);
            Block &s = blk[(size_t)(g * per_gpu + ((b / n_gpus) % per_gpu))];
            {
                std::unique_lock<std::mutex> lk(s.m);
                s.cv.wait(lk, [&] { return s.ready || failed.load(); });
                if (!s.ready) { error = s.error.empty() ? "a worker failed" : s.error; break; }
                s.ready = false;
            }
            const size_t cnt = (size_t)(s.hi - s.lo) * rings;
            bool bad = false;                                   // checkRange (:362-366)
            for (size_t i = 0; i < cnt && !bad; i++) bad = !std::isfinite(s.host[i]);
            if (bad) {
                printf("\nDist contains NaN\n");
                fail_all("NaN");
                for (auto &w : workers) w.join();
                return -1;
            }
            out.append_rows(hd, s.host, (uint64_t)(s.hi - s.lo), io_threads);
            printf("\rStep: %lld / %lld", (long long)s.hi, (long long)n_pairs);
            fflush(stdout);
            std::lock_guard<std::mutex> lk(s.m);
            s.free_ = true;
            s.cv.notify_all();
        }
        if (!error.empty()) fail_all(error);
        for (auto &w : workers) w.join();
        if (!error.empty()) throw AbiError("dlco_fulldist_build_host: " + error);
        const double t_compute = now_sec() - t0;
        out.close();
        for (auto &b : blk) cudaFreeHost(b.host);
        for (int g = 0; g < n_gpus; g++) dlco_bank_destroy(ctxs[g]->h, banks[g]);
        printf("\nDone.\n\n");
        printf("Total: %.09f sec\n\n", now_sec() - t0);
        fprintf(stderr, "[dlco_b200] %lld pairs, %.1f pairs/s including HDF5 output (%d GPU%s)\n", (long long)n_pairs,
                n_pairs / std::max(t_compute, 1e-9), n_gpus, n_gpus > 1 ? "s" : "");
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\ncomp-fulldists: %s\n", e.what());
        return 2;
    }
}
