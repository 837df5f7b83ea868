// h5tool -- inspect / re-write HDF5 files through h5mini (used by tests/test_h5mini.py to check the
// C++ reader against the reference's committed .h5 files and the writer against an independent
// reader).
//   h5tool ls   <file>                      one line per dataset: name dtype dims [chunk] [deflate]
//   h5tool raw  <file> <dataset> <out.bin> [lo hi]   rows [lo, hi) as raw little-endian bytes
//   h5tool copy <in> <out> [chunk_rows] [deflate] [threads]   every dataset, through Reader -> Writer
//   h5tool pack <out> {<name> <dtype> <ndims> <dims..> <file.bin> <chunk_rows|0> <deflate>}...
//   h5tool gen  <out> <rows> <chunk_rows> <deflate>   formula data in the pipeline's dataset shapes
//   h5tool append-test <out>                the pr-learn `w` pattern: unlimited rows, flush per row
#include "h5mini.hpp"
#include <cstdlib>

using namespace h5mini;

static const char *tname(DType t) { return t == U8 ? "u8" : t == I32 ? "i32" : t == F32 ? "f32" : "f64"; }

int main(int argc, char **argv)
{
    try {
        if (argc >= 3 && !strcmp(argv[1], "ls")) {
            Reader r(argv[2]);
            for (auto &n : r.names()) {
                const DatasetInfo &d = r.info(n);
                printf("%s %s", n.c_str(), tname(d.dtype));
                for (auto x : d.dims) printf(" %llu", (unsigned long long)x);
                if (d.layout == 2) {
                    printf(" chunk");
                    for (auto x : d.chunk) printf(" %u", x);
                }
                if (d.deflate) printf(" deflate");
                if (!d.maxdims.empty() && d.maxdims[0] == UNDEF) printf(" unlimited");
                printf("\n");
            }
            return 0;
        }
        if (argc >= 5 && !strcmp(argv[1], "raw")) {
            Reader r(argv[2]);
            const DatasetInfo &d = r.info(argv[3]);
            uint64_t lo = 0, hi = d.rows();
            if (argc >= 7) { lo = strtoull(argv[5], 0, 10); hi = strtoull(argv[6], 0, 10); }
            std::vector<uint8_t> buf((hi - lo) * d.row_elems() * dtype_size(d.dtype));
            r.read_rows(argv[3], lo, hi, buf.data(), 4);
            FILE *f = fopen(argv[4], "wb");
            if (!f || fwrite(buf.data(), 1, buf.size(), f) != buf.size()) throw Error("cannot write output");
            fclose(f);
            return 0;
        }
        if (argc >= 4 && !strcmp(argv[1], "copy")) {
            const uint32_t chunk_rows = argc >= 5 ? (uint32_t)atoi(argv[4]) : 0;
            const int level = argc >= 6 ? atoi(argv[5]) : 0, threads = argc >= 7 ? atoi(argv[6]) : 1;
            Reader r(argv[2]);
            Writer w(argv[3]);
            for (auto &n : r.names()) {
                const DatasetInfo &d = r.info(n);
                std::vector<uint8_t> buf(d.rows() * d.row_elems() * dtype_size(d.dtype));
                r.read_rows(n, 0, d.rows(), buf.data(), threads);
                if (!chunk_rows) {
                    w.write_contiguous(n, d.dtype, d.dims, buf.data());
                } else {
                    std::vector<uint64_t> trailing(d.dims.begin() + 1, d.dims.end());
                    const int h = w.create_chunked(n, d.dtype, trailing, chunk_rows, level, d.maxdims[0] == UNDEF);
                    // feed in uneven pieces to exercise the partial-chunk buffering
                    const uint64_t rb = d.row_elems() * dtype_size(d.dtype);
                    uint64_t row = 0, step = 1;
                    while (row < d.rows()) {
                        const uint64_t n2 = std::min<uint64_t>(step, d.rows() - row);
                        w.append_rows(h, buf.data() + row * rb, n2, threads);
                        row += n2;
                        step = step * 3 + 1;
                    }
                }
            }
            w.close();
            return 0;
        }
        if (argc >= 3 && !strcmp(argv[1], "append-test")) {
            Writer w(argv[2]);
            const int h = w.create_chunked("w", F32, {5120}, 1, 9, true);
            std::vector<float> row(5120);
            for (int k = 0; k < 5; k++) {
                for (int i = 0; i < 5120; i++) row[i] = (i % 97 == k) ? (float)(k + 1) * 0.25f : 0.0f;
                w.append_rows(h, row.data(), 1);
                w.flush();                      // the file is complete after every accepted model
            }
            w.close();
            return 0;
        }
        if (argc >= 3 && !strcmp(argv[1], "pack")) {
            // pack <out> {<name> <u8|i32|f32|f64> <ndims> <d0> .. <file.bin> <chunk_rows|0> <deflate>}...
            Writer w(argv[2]);
            int i = 3;
            while (i + 2 < argc) {
                const std::string name = argv[i++], tn = argv[i++];
                const DType t = tn == "u8" ? U8 : tn == "i32" ? I32 : tn == "f64" ? F64 : F32;
                const int nd = atoi(argv[i++]);
                if (nd < 1 || nd > 4 || i + nd + 3 > argc) throw Error("pack: bad dataset spec for " + name);
                std::vector<uint64_t> dims;
                for (int k = 0; k < nd; k++) dims.push_back(strtoull(argv[i++], 0, 10));
                const char *bin = argv[i++];
                const uint32_t cr = (uint32_t)atoi(argv[i++]);
                const int level = atoi(argv[i++]);
                uint64_t bytes = dtype_size(t);
                for (auto d : dims) bytes *= d;
                std::vector<uint8_t> buf(bytes);
                FILE *f = fopen(bin, "rb");
                if (!f || fread(buf.data(), 1, bytes, f) != bytes) throw Error(std::string("pack: cannot read ") + bin);
                fclose(f);
                if (!cr) {
                    w.write_contiguous(name, t, dims, buf.data());
                } else {
                    std::vector<uint64_t> trailing(dims.begin() + 1, dims.end());
                    const int h = w.create_chunked(name, t, trailing, cr, level, false);
                    w.append_rows(h, buf.data(), dims[0], 4);
                }
            }
            w.close();
            return 0;
        }
        if (argc >= 6 && !strcmp(argv[1], "gen")) {
            // gen <out> <rows> <chunk_rows> <deflate>: the dataset shapes of the DLCO pipeline with
            // formula data (tests/test_h5mini.py recomputes them): Distance f32 rows x 40,
            // Label u8 rows x 1, Indices i32 rows x 4, Patches u8 (rows/8) x 64 x 64, RingParams f64 7 x 3
            const uint64_t rows = strtoull(argv[3], 0, 10);
            const uint32_t cr = (uint32_t)atoi(argv[4]);
            const int level = atoi(argv[5]);
            Writer w(argv[2]);
            const int hd = w.create_chunked("Distance", F32, {40}, cr, level, false);
            const int hl = w.create_chunked("Label", U8, {1}, cr, level, false);
            const int hi = w.create_chunked("Indices", I32, {4}, cr, 0, false);
            const int hp = w.create_chunked("Patches", U8, {64, 64}, cr, level, false);
            std::vector<float> d(40);
            for (uint64_t r = 0; r < rows; r++) {
                for (int c = 0; c < 40; c++) d[c] = (float)r * 0.5f + (float)c * 0.001f;
                const uint8_t lab = (uint8_t)(r % 3 == 0);
                const int32_t idx[4] = {(int32_t)r, (int32_t)(r / 2), (int32_t)(r + 7), (int32_t)(r / 3)};
                w.append_rows(hd, d.data(), 1, 2);
                w.append_rows(hl, &lab, 1, 2);
                w.append_rows(hi, idx, 1, 2);
            }
            std::vector<uint8_t> patch(4096);
            for (uint64_t p = 0; p < rows / 8; p++) {
                for (int i = 0; i < 4096; i++) patch[i] = (uint8_t)((p * 31 + i * 7 + (i >> 6)) & 255);
                w.append_rows(hp, patch.data(), 1, 2);
            }
            double ring[21];
            for (int i = 0; i < 21; i++) ring[i] = 0.125 * i - 1.0 / 3.0;
            w.write_contiguous("RingParams", F64, {7, 3}, ring);
            w.close();
            return 0;
        }
        fprintf(stderr, "usage: h5tool ls|raw|copy|pack|append-test|gen ...\n");
        return 1;
    } catch (const std::exception &e) {
        fprintf(stderr, "h5tool: %s\n", e.what());
        return 2;
    }
}
