// conv-impatches -- drop-in for the reference tool (src/conv-impatches.cpp): the Brown / Winder patch data set
// (patchesNNNN.bmp: 1024 x 1024 bitmaps of 16 x 16 patches; m50_500000_500000_0.txt: the pair list) ->
// "Indices" i32 N x 4 (patch1, point1, patch2, point2) and "Patches" u8 P x 64 x 64, the input of
// comp-fulldists / comp-uprjdists.  Host only.  The reference decodes the bitmaps with cv::imread(GRAYSCALE);
// this reads the BMP directly (8-bit palettised -- the data set's format -- and 24 / 32-bit, bottom-up or
// top-down, uncompressed) with OpenCV's fixed-point grey conversion for coloured palettes.
//
//   conv-impatches patch_dir_path dst_h5_file
#include "common.hpp"

using namespace dlco;

static inline uint8_t to_grey(unsigned b, unsigned g, unsigned r)
{
    return (uint8_t)((b * 1868u + g * 9617u + r * 4899u + (1u << 13)) >> 14);      // cv: B2Y, G2Y, R2Y, shift 14
}

// -> rows x cols grey image, false when the file does not exist
static bool read_bmp_grey(const std::string &path, std::vector<uint8_t> &img, int &rows, int &cols)
{
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    std::vector<uint8_t> buf;
    uint8_t tmp[65536];
    size_t n;
    while ((n = fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
    fclose(f);
    auto u16 = [&](size_t o) { return (unsigned)buf.at(o) | (unsigned)buf.at(o + 1) << 8; };
    auto u32 = [&](size_t o) { return (uint32_t)u16(o) | (uint32_t)u16(o + 2) << 16; };
    if (buf.size() < 54 || buf[0] != 'B' || buf[1] != 'M') throw h5mini::Error("not a BMP file: " + path);
    const uint32_t data_off = u32(10), hdr = u32(14);
    if (hdr < 40) throw h5mini::Error("unsupported BMP header: " + path);
    const int32_t w = (int32_t)u32(18), h_raw = (int32_t)u32(22);
    const unsigned bpp = u16(28), comp = u32(30);
    uint32_t ncol = u32(46);
    if (comp != 0 || (bpp != 8 && bpp != 24 && bpp != 32) || w <= 0 || h_raw == 0)
        throw h5mini::Error("unsupported BMP (need uncompressed 8/24/32 bpp): " + path);
    const bool top_down = h_raw < 0;
    const int h = top_down ? -h_raw : h_raw;
    uint8_t pal[256];
    if (bpp == 8) {
        if (ncol == 0 || ncol > 256) ncol = 256;
        for (uint32_t i = 0; i < 256; i++) pal[i] = 0;
        for (uint32_t i = 0; i < ncol; i++) {
            const size_t o = 14 + hdr + 4 * (size_t)i;
            pal[i] = to_grey(buf.at(o), buf.at(o + 1), buf.at(o + 2));
        }
    }
    const size_t stride = ((size_t)w * bpp / 8 + 3) & ~(size_t)3;
    if ((size_t)data_off + stride * h > buf.size()) throw h5mini::Error("truncated BMP: " + path);
    rows = h;
    cols = w;
    img.resize((size_t)h * w);
    for (int y = 0; y < h; y++) {
        const uint8_t *src = &buf[data_off + stride * (size_t)(top_down ? y : h - 1 - y)];
        uint8_t *dst = &img[(size_t)y * w];
        if (bpp == 8)
            for (int x = 0; x < w; x++) dst[x] = pal[src[x]];
        else
            for (int x = 0; x < w; x++) dst[x] = to_grey(src[x * (bpp / 8)], src[x * (bpp / 8) + 1], src[x * (bpp / 8) + 2]);
    }
    return true;
}

int main(int argc, char **argv)
{
    const int sChunk = 256;
    bool help = false;
    const char *dir = nullptr, *dst = nullptr;
    for (int i = 1; i < argc; i++) {
        if (argv[i][0] == '-') {
            if (!strcmp(argv[i], "-help")) { help = true; continue; }
            printf("ERROR: Invalid %s option.\n", argv[i]);
            help = true;
        } else if (!dir) dir = argv[i];
        else if (!dst) dst = argv[i];
    }
    if (!dir || !dst) help = true;
    if (help) {
        printf("\nUsage: conv-imgpatches patch_dir_path dst_h5_file\n\n");
        return 1;
    }
    try {
        const int io_threads = host_threads();
        const std::string mname = std::string(dir) + "/m50_500000_500000_0.txt";
        printf("Open index file: %s\n", mname.c_str());
        FILE *mf = fopen(mname.c_str(), "r");
        if (!mf) {
            printf("ERROR: File m50_500000_500000_0.txt not found.\n\n");
            return 1;
        }
        // "patchID1 3DpointID1 unused1 patchID2 3DpointID2 unused2" per line (src/conv-impatches.cpp:146-148)
        std::vector<int32_t> indices;
        int num_patches = 0;
        char line[4096];
        while (fgets(line, sizeof(line), mf)) {
            int p1 = 0, q1 = 0, p2 = 0, q2 = 0;
            sscanf(line, "%i %i %*s %i %i %*s", &p1, &q1, &p2, &q2);
            indices.insert(indices.end(), {p1, q1, p2, q2});
            num_patches = std::max(num_patches, std::max(p1, p2));
        }
        fclose(mf);
        const int64_t n_lines = (int64_t)indices.size() / 4;
        printf("Export Indices: #%lld\n", (long long)n_lines);
        h5mini::Writer out(dst);
        {
            const int h = out.create_chunked("Indices", h5mini::I32, {4}, sChunk, 0, false);     // H5_NONE (:129)
            Progress bar;
            out.append_rows(h, indices.data(), (uint64_t)n_lines, io_threads);
            bar.update(1.0);
        }
        num_patches += 1;                                            // ids start at 0
        printf("Export Patches: #%d\n", num_patches);
        const int hp = out.create_chunked("Patches", h5mini::U8, {64, 64}, sChunk, deflate_level(), false);
        std::vector<uint8_t> img, block((size_t)256 * 4096);
        Progress bar;
        int idx = 0;
        for (int i = 0; i <= num_patches / 256; i++) {
            char name[64];
            snprintf(name, sizeof(name), "/patches%04i.bmp", i);
            int rows = 0, cols = 0;
            int chunk = 0;
            const bool have = idx < num_patches && read_bmp_grey(std::string(dir) + name, img, rows, cols);
            if (idx < num_patches && !have) throw h5mini::Error(std::string("missing bitmap ") + dir + name);
            if (have && (rows < 1024 || cols < 1024)) throw h5mini::Error(std::string("bitmap smaller than 1024 x 1024: ") + dir + name);
            for (int r = 0; r < 16 && idx < num_patches; r++)
                for (int c = 0; c < 16 && idx < num_patches; c++) {
                    for (int y = 0; y < 64; y++)
                        memcpy(&block[(size_t)chunk * 4096 + y * 64], &img[(size_t)(r * 64 + y) * cols + c * 64], 64);
                    idx++;
                    chunk++;
                }
            if (chunk) out.append_rows(hp, block.data(), (uint64_t)chunk, io_threads);
            bar.update((double)i / (double)std::max(1, num_patches / 256));
        }
        bar.update(1.0);
        out.close();
        printf("Done.\n\n");
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "\nconv-impatches: %s\n", e.what());
        return 2;
    }
}
