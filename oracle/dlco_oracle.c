/* C restatement of the heavy CPU pieces of the DLCO hot path -- TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product never links or dlopens it.
 *
 *  - dlco_oracle_pr_bank      : src/gen-poolregion.cpp:51-80,241-275 (bit-exact bank;
 *                               pinned against workspace/opencv/vgg_generated_*.i)
 *  - dlco_oracle_pool_pair    : src/comp-fulldists.cpp:328-342 for one pair (double acc)
 *  - dlco_oracle_hinge_pairsum: src/kernelop-opencv.cu:49-66 + cuda::sum, brute force
 *
 * The reference itself is not buildable here (needs OpenCV C++ with cudaarithm + hdf,
 * libhdf5, OpenBLAS headers), so this is a "port", not oracle/_ref.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* src/gen-poolregion.cpp:51-80.  out: PatchSize*PatchSize floats, row index = x. */
static void pr_filter(int ps, double x0, double y0, double sigma, float *out, float *ext)
{
    int rExt = (int)ceil(3.0f * sigma);
    int n = ps + 2 * rExt;
    memset(ext, 0, sizeof(float) * (size_t)n * n);
    x0 = x0 + 0.5f * (1.0f + ps) + rExt;
    y0 = y0 + 0.5f * (1.0f + ps) + rExt;
    for (int y = (int)floor(y0 - 3.0f * sigma); y <= (int)ceil(y0 + 3.0f * sigma); y++)
        for (int x = (int)floor(x0 - 3.0f * sigma); x <= (int)ceil(x0 + 3.0f * sigma); x++) {
            double dx = x - x0, dy = y - y0;
            double r2 = dx * dx + dy * dy;
            ext[(size_t)(x - 1) * n + (y - 1)] = (float)exp(-r2 / (2 * sigma * sigma));
        }
    double s = 0.0;
    for (int i = 0; i < ps; i++)
        for (int j = 0; j < ps; j++) {
            float v = ext[(size_t)(i + rExt) * n + (j + rExt)];
            out[i * ps + j] = v;
            s += v;
        }
    float inv = (float)(1.0 / s);          /* cv::Mat /= s  ==  *= float(1/s) */
    for (int i = 0; i < ps * ps; i++) out[i] *= inv;
}

/* rings: nRings x 3 doubles (r, phi, sigma); filters: nRings*8 x ps*ps; params: nRings*8 x 3 */
void dlco_oracle_pr_bank(int ps, int nRings, const double *rings, float *filters, float *params)
{
#pragma omp parallel
    {
        float *ext = (float *)malloc(sizeof(float) * (size_t)(ps + 2 * 64) * (ps + 2 * 64));
#pragma omp for schedule(dynamic, 8)
        for (int c = 0; c < nRings; c++) {
            double r = rings[3 * c], phi = rings[3 * c + 1], sigma = rings[3 * c + 2];
            double xc = r * cos(phi), yc = r * sin(phi);
            double off[8][2] = {{yc, xc}, {yc, -xc}, {-yc, xc}, {-yc, -xc},
                                {xc, yc}, {xc, -yc}, {-xc, -yc}, {-xc, yc}};
            for (int i = 0; i < 8; i++) {
                size_t k = (size_t)c * 8 + i;
                pr_filter(ps, off[i][1], off[i][0], sigma, filters + k * ps * ps, ext);
                if (params) {
                    params[3 * k + 0] = (float)off[i][1];
                    params[3 * k + 1] = (float)off[i][0];
                    params[3 * k + 2] = (float)sigma;
                }
            }
        }
        free(ext);
    }
}

/* One pair, double accumulation: dist[ring] = sum_{j<8} sum_{b<8}
 * (min(F[8r+j].T2[:,b],1) - min(F[8r+j].T1[:,b],1))^2.   t1/t2: K x 8 row-major. */
void dlco_oracle_pool_pair(int nF, int K, const float *filters, const float *t1,
                           const float *t2, double *dist)
{
#pragma omp parallel for schedule(static)
    for (int ring = 0; ring < nF / 8; ring++) {
        double acc = 0.0;
        for (int j = 0; j < 8; j++) {
            const float *f = filters + (size_t)(ring * 8 + j) * K;
            double p1[8] = {0}, p2[8] = {0};
            for (int k = 0; k < K; k++) {
                double fv = f[k];
                if (fv == 0.0) continue;
                for (int b = 0; b < 8; b++) {
                    p1[b] += fv * t1[k * 8 + b];
                    p2[b] += fv * t2[k * 8 + b];
                }
            }
            for (int b = 0; b < 8; b++) {
                double a = p1[b] < 1.0 ? p1[b] : 1.0, c = p2[b] < 1.0 ? p2[b] : 1.0;
                acc += (c - a) * (c - a);
            }
        }
        dist[ring] = acc;
    }
}

double dlco_oracle_hinge_pairsum(const float *pos, int n1, const float *neg, int n2, double margin)
{
    double tot = 0.0;
#pragma omp parallel for reduction(+ : tot)
    for (int i = 0; i < n1; i++) {
        double p = (double)pos[i] + margin, s = 0.0;
        for (int j = 0; j < n2; j++) {
            double d = p - (double)neg[j];
            s += d > 0.0 ? d : 0.0;
        }
        tot += s;
    }
    return tot;
}

/* pr-learn RDA chain, fp32 "reference order" (src/pr-learn.cpp:303-329, single-thread draw
 * order): same arithmetic as dlco_oracle.py:pr_rda_steps(dt=f32).  rng_state: cv::RNG state
 * (multiply-with-carry, uniform(0,n) = next() % n), updated.  The CPU baseline of row c1. */
void dlco_oracle_pr_rda(const float *dists, long ld, int F, const int *idx_pos, int n_pos_trn,
                        const int *idx_neg, int n_neg_trn, float mu, float gamma,
                        unsigned long long t0, unsigned long long n_steps,
                        unsigned long long *rng_state, float *w, float *df_avg)
{
    unsigned long long s = *rng_state;
    float *diff = (float *)malloc(sizeof(float) * (size_t)F);
    for (unsigned long long k = 0; k < n_steps; k++) {
        const unsigned long long t = t0 + k;
        s = (unsigned long long)(unsigned)s * 4164903690ULL + (s >> 32);
        const int ip = (int)((unsigned)s % (unsigned)n_pos_trn);
        s = (unsigned long long)(unsigned)s * 4164903690ULL + (s >> 32);
        const int in = (int)((unsigned)s % (unsigned)n_neg_trn);
        const float *a = dists + (size_t)idx_pos[ip] * ld, *b = dists + (size_t)idx_neg[in] * ld;
        double acc = 0.0;
        for (int j = 0; j < F; j++) {
            diff[j] = a[j] - b[j];
            acc += (double)w[j] * (double)diff[j];
        }
        const float f = (float)acc;
        const float decay = (float)((double)t * (1.0 / (double)(t + 1)));
        const float inv = (float)(1.0 / (double)(t + 1));
        const double alpha = -sqrt((double)(t + 1)) / (double)gamma;
        const float fa = (float)alpha, fb = (float)(alpha * (double)mu);
        const int active = f > -1.0f;
        for (int j = 0; j < F; j++) {
            float d = df_avg[j] * decay;
            if (active) d = diff[j] * inv + d;
            df_avg[j] = d;
            const float v = d * fa + fb;
            w[j] = v > 0.0f ? v : 0.0f;
        }
    }
    free(diff);
    *rng_state = s;
}
