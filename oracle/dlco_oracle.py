"""CPU restatement of the DLCO training hot path -- TEST INFRASTRUCTURE ONLY.

This module is the parity *checker*.  Only `tests/`, `tests/golden/make_golden.py`,
`__graft_entry__.smoke()` and `bench.py`'s `cpu_baseline` / `--impl reference` legs
may import it; the product (`opencv-dlco_b200/`) never does and fails loudly when
its CUDA library is missing.

Every function cites the reference file:line it restates (paths relative to
`/root/reference/`).  The reference's arithmetic partly lives in un-vendored
third-party code (OpenCV >= 3.1, un-pinned; OpenBLAS): those primitives are
restated from their documented behaviour and pinned two ways, see
`tests/test_oracle_golden.py`:
  * against the reference tree's own committed outputs (`workspace/opencv/*.i`,
    `workspace/pr-learn/*.h5`, `workspace/pj-learn/*.h5` and their logs), through
    the fixtures in `tests/golden/` made by `tests/golden/make_golden.py`;
  * against the same OpenCV primitives in `cv2` 4.13 (GaussianBlur, filter2D,
    magnitude, sort, randShuffle, contourArea) when `cv2` is importable.
The image-dependent stages (get_desc -> distances -> learners) have no golden
outputs in the reference tree (the Brown dataset is absent): their parity is
pinned only through those primitives, i.e. "parity unpinned" above them.

dtype convention: `dt=np.float32` mimics the reference's fp32 evaluation order as
far as it is defined; `dt=np.float64` is the "truth" mode used for tolerances.
"""
import math

import numpy as np

CV_PI = 3.1415926535897932384626433832795
f32 = np.float32


# ---------------------------------------------------------------------------
# cv::RNG (multiply-with-carry) -- used by src/pr-learn.cpp:241-243,306-310 and
# src/pj-learn.cpp:225-227,311-316.  OpenCV core: state = (u32)state*4164903690
# + (state >> 32); RNG() state 0xffffffff; uniform(int a,int b) = next()%(b-a)+a.
# ---------------------------------------------------------------------------
class CvRNG:
    COEFF = 4164903690

    def __init__(self, seed=0xFFFFFFFF):
        self.state = seed if seed != 0 else 0xFFFFFFFF

    def next(self):
        s = self.state
        s = ((s & 0xFFFFFFFF) * self.COEFF + (s >> 32)) & 0xFFFFFFFFFFFFFFFF
        self.state = s
        return s & 0xFFFFFFFF

    def uniform(self, a, b):
        return a if a == b else a + self.next() % (b - a)

    def draws(self, n):
        """n raw 32-bit outputs as a numpy array (vectorised where possible)."""
        out = np.empty(n, np.uint64)
        s = self.state
        for i in range(n):
            s = ((s & 0xFFFFFFFF) * self.COEFF + (s >> 32)) & 0xFFFFFFFFFFFFFFFF
            out[i] = s & 0xFFFFFFFF
        self.state = s
        return out


def rand_shuffle(arr, rng):
    """cv::randShuffle(vector<int>) with iterFactor=1: for i: swap(a[i], a[next()%n]).
    The reference calls it on the *global* theRNG() (state 0xffffffff in a fresh
    process), IdxPos first then IdxNeg (src/pr-learn.cpp:242-243)."""
    a = list(arr)
    n = len(a)
    for i in range(n):
        j = rng.next() % n
        a[i], a[j] = a[j], a[i]
    return a


def split_train_val(labels, n_div=0.80):
    """src/pr-learn.cpp:230-254 == src/pj-learn.cpp:214-237.
    -> IdxPos, IdxNeg (shuffled), nPosTrn, nNegTrn."""
    labels = np.asarray(labels).reshape(-1)
    idx_pos = [int(i) for i in np.nonzero(labels == 1)[0]]
    idx_neg = [int(i) for i in np.nonzero(labels == 0)[0]]
    g = CvRNG()  # theRNG() default state
    idx_pos = rand_shuffle(idx_pos, g)
    idx_neg = rand_shuffle(idx_neg, g)
    # size_t nPosTrn = IdxPos.size() * nDiv;  (float nDiv=0.80 -> double product, trunc)
    n_pos_trn = int(len(idx_pos) * float(f32(n_div)))
    n_neg_trn = int(len(idx_neg) * float(f32(n_div)))
    return np.array(idx_pos, np.int64), np.array(idx_neg, np.int64), n_pos_trn, n_neg_trn


# ---------------------------------------------------------------------------
# gen-poolregion: src/gen-poolregion.cpp:51-80 (filter), :163-213 (ring grid),
# :241-295 (8 symmetric members per ring).
# ---------------------------------------------------------------------------
def ring_params(patch_size=64, nr=32, nphi=5, nsigma=32):
    """src/gen-poolregion.cpp:163-172,189-213; loops accumulate in double as written."""
    patch_rad = math.floor((patch_size - 1) / 2)
    r0, r1 = 0.0, float(patch_rad)
    phi0, phi1 = 0.0, CV_PI / 4.0
    sigma0, sigma1 = 0.5, math.ceil(patch_rad / 2.0)
    out = []
    s = sigma0
    while s <= sigma1:
        p = phi0
        while p <= phi1:
            r = r0
            while r <= r1:
                out.append((r, p, s))
                r = r + ((r1 - r0) / (nr - 1))
            p = p + ((phi1 - phi0) / (nphi - 1))
        s = s + ((sigma1 - sigma0) / (nsigma - 1))
    return np.array(out, np.float64)


def ring_offsets(r, phi):
    """src/gen-poolregion.cpp:247-257 -> 8 rows (Offsets[:,0], Offsets[:,1])."""
    xc = r * math.cos(phi)
    yc = r * math.sin(phi)
    return [(yc, xc), (yc, -xc), (-yc, xc), (-yc, -xc),
            (xc, yc), (xc, -yc), (-xc, -yc), (-xc, yc)]


def get_pr_filter(patch_size, x0, y0, sigma):
    """src/gen-poolregion.cpp:51-80, bit-exact: f64 math, exp rounded to f32 on store,
    row index = x; unit-sum normalisation is OpenCV's `Mat /= s`, which multiplies by
    float(1/s) (checked against workspace/opencv/*.i)."""
    r_ext = math.ceil(3.0 * sigma)
    n = patch_size + 2 * r_ext
    pr = np.zeros((n, n), f32)
    x0 = x0 + float(f32(0.5) * (f32(1.0) + f32(patch_size))) + r_ext
    y0 = y0 + float(f32(0.5) * (f32(1.0) + f32(patch_size))) + r_ext
    two_s2 = 2 * sigma * sigma
    for y in range(math.floor(y0 - 3.0 * sigma), math.ceil(y0 + 3.0 * sigma) + 1):
        dy = y - y0
        for x in range(math.floor(x0 - 3.0 * sigma), math.ceil(x0 + 3.0 * sigma) + 1):
            dx = x - x0
            r2 = dx * dx + dy * dy
            pr[x - 1, y - 1] = f32(math.exp(-r2 / two_s2))
    pr = pr[r_ext:r_ext + patch_size, r_ext:r_ext + patch_size].copy()
    s = float(np.sum(pr.astype(np.float64)))
    return pr * f32(1.0 / s)


def pr_filters_for_rings(rings, patch_size=64):
    """Filters (8 per ring, reference order) and PRParams rows for `rings` (k x 3).
    src/gen-poolregion.cpp:241-275."""
    filt = np.zeros((len(rings) * 8, patch_size * patch_size), f32)
    prp = np.zeros((len(rings) * 8, 3), f32)
    k = 0
    for (r, phi, sigma) in rings:
        for (o0, o1) in ring_offsets(r, phi):
            filt[k] = get_pr_filter(patch_size, o1, o0, sigma).reshape(-1)
            prp[k] = (f32(o1), f32(o0), f32(sigma))
            k += 1
    return filt, prp


def pr_params(rings):
    """PRParams only (40960 x 3 f32), src/gen-poolregion.cpp:268-271."""
    prp = np.zeros((len(rings) * 8, 3), f32)
    k = 0
    for (r, phi, sigma) in rings:
        for (o0, o1) in ring_offsets(r, phi):
            prp[k] = (f32(o1), f32(o0), f32(sigma))
            k += 1
    return prp


# ---------------------------------------------------------------------------
# SelectPRFilters: src/misc.cpp:78-169
# ---------------------------------------------------------------------------
def select_pr_filters(pr_filters, w):
    """rows 8i+j with w[i]>0 and any non-zero -> first-occurrence unique -> sortrows."""
    w = np.asarray(w).reshape(-1)
    assert w.size * 8 == pr_filters.shape[0]
    rows = []
    for i in np.nonzero(w > 0)[0]:
        for j in range(8):
            row = pr_filters[8 * i + j]
            if np.count_nonzero(row):
                rows.append(row)
    uniq = []
    for row in rows:
        if not any(np.array_equal(row, u) for u in uniq):
            uniq.append(row)
    uniq.sort(key=lambda v: v.tolist())     # lexicographic == misc.cpp:124-166
    return np.array(uniq, f32)


# ---------------------------------------------------------------------------
# get_desc: src/vgg-desc.cpp:41-152
# ---------------------------------------------------------------------------
def gaussian_kernel_1d(sigma, dt=f32):
    """cv::GaussianBlur(ksize=(0,0), sigma) on a CV_32F image: ksize =
    cvRound(sigma*4*2+1)|1 (=13 for 1.4); cv::getGaussianKernel computes
    exp(-x^2/(2 sigma^2)) in double, normalises, stores as float."""
    ksize = int(round(sigma * 4 * 2 + 1)) | 1
    x = np.arange(ksize, dtype=np.float64) - (ksize - 1) * 0.5
    k = np.exp(-(x * x) / (2.0 * sigma * sigma))
    k = k / k.sum()
    return k.astype(dt)


def _fma32(a, b, c):
    """float32 fused multiply-add, elementwise: the product of two f32 is exact in f64, so
    f32(f64(a)*f64(b) + f64(c)) is the fused result (up to a double rounding that needs a 29-bit
    coincidence)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def _sep_blur_replicate(img, k):
    """cv::GaussianBlur -> cv::sepFilter2D, BORDER_REPLICATE, rows then columns.
    fp32: OpenCV's own evaluation order, established against cv2 4.13 (AVX2 dispatch) bit for bit
    (tests/test_oracle_golden.py): the row filter accumulates the taps left to right with fused
    multiply-adds (RowVec_32f), the column filter is the symmetric form
    k[c]*x0 + sum_j k[c+j]*(x[+j] + x[-j]) with fused multiply-adds (SymmColumnVec_32f).
    fp64: plain sums (truth mode)."""
    r = len(k) // 2
    dt = img.dtype
    w = img.shape[1]
    h = img.shape[0]
    p = np.pad(img, ((0, 0), (r, r)), mode="edge")
    if dt == f32:
        tmp = np.zeros_like(img)
        for j in range(len(k)):
            tmp = _fma32(p[:, j:j + w], k[j], tmp)
        p = np.pad(tmp, ((r, r), (0, 0)), mode="edge")
        out = (k[r] * p[r:r + h, :]).astype(f32)
        for j in range(1, r + 1):
            out = _fma32((p[r + j:r + j + h, :] + p[r - j:r - j + h, :]).astype(f32), k[r + j], out)
        return out
    tmp = np.zeros_like(img)
    for j in range(len(k)):
        tmp = tmp + k[j] * p[:, j:j + w]
    p = np.pad(tmp, ((r, r), (0, 0)), mode="edge")
    out = np.zeros_like(img)
    for j in range(len(k)):
        out = out + k[j] * p[j:j + h, :]
    return out.astype(dt)


def quantile_threshold(gmag, q=0.8):
    """src/vgg-desc.cpp:104-129: scipy-mquantiles rule (alphap=betap=0.5) on the
    ascending sort of all magnitudes; the index arithmetic is float32 as written; T is two
    products and a sum, each rounded (the reference is built without -march, no contraction)."""
    s = np.sort(gmag.reshape(-1))
    n = s.size
    aleph = f32(f32(n) * f32(q)) + f32(0.5)
    k = int(math.floor(float(aleph)))
    if k >= n - 1:
        k = n - 1
    if k <= 1:
        k = 1
    gamma = f32(aleph - f32(k))
    gamma = min(max(gamma, f32(0.0)), f32(1.0))
    dt = gmag.dtype.type
    return dt(dt(dt(1.0) - dt(gamma)) * s[k - 1]) + dt(dt(gamma) * s[k])


def get_desc_parts(patch, n_bins=8, init_sigma=1.4, b_norm=True, dt=f32, atan2_float=False):
    """src/vgg-desc.cpp:41-132 -> (w1*GMag, w2*GMag, Bin1, Bin2, T) as 64x64 images (row = y).
    dt=f32 follows OpenCV's arithmetic to the bit (pinned by tests/golden/desc_kat.npz, generated
    by driving cv2 through the reference's call sequence, tests/golden/make_desc_golden.py):
      * blur: see _sep_blur_replicate;  gradients: exact differences;
      * cv::magnitude = sqrt(fma(x, x, y*y)) (the non-IPP kernel; IPP's is +-2 ulp around it);
      * GAngle = float(atan2(double Iy, double Ix) + CV_PI) (atan2_float: float(double(atan2f) + CV_PI));
      * GAngle / AngleStep - 0.5f and GMag /= s are scaled convertTo's: x * float(1/s) + b, fused."""
    dt = np.dtype(dt).type
    img = np.asarray(patch).astype(dt)
    # `float InitSigma` (:41) is promoted to GaussianBlur's double sigma: 1.4f = 1.39999997615..., not 1.4
    img = _sep_blur_replicate(img, gaussian_kernel_1d(float(f32(init_sigma)), dt))       # :46
    px = np.pad(img, ((0, 0), (1, 1)), mode="edge")
    py = np.pad(img, ((1, 1), (0, 0)), mode="edge")
    ix = px[:, 2:] - px[:, :-2]                                              # :50-53
    iy = py[2:, :] - py[:-2, :]
    if dt is f32:
        gmag = np.sqrt(_fma32(ix, ix, (iy * iy).astype(f32))).astype(f32)    # :59
        if atan2_float:
            gang = (np.arctan2(iy, ix).astype(np.float64) + CV_PI).astype(f32)
        else:
            gang = (np.arctan2(iy.astype(np.float64), ix.astype(np.float64)) + CV_PI).astype(f32)   # :64-66
        step = f32(2.0 * CV_PI / float(f32(n_bins)))                         # :69
        ratio = _fma32(gang, f32(1.0 / float(step)), f32(-0.5))              # :70
    else:
        gmag = np.sqrt(ix * ix + iy * iy)
        gang = np.arctan2(iy, ix) + CV_PI
        ratio = gang / dt(2.0 * CV_PI / n_bins) - dt(0.5)
    off1 = (ratio - np.floor(ratio)).astype(dt)                              # :72-74
    b1 = np.ceil(ratio - dt(1.0))                                            # :84-91
    b1 = np.where(b1 == -1.0, n_bins - 1, b1).astype(np.int64)
    b2 = np.where(b1 + 1 > n_bins - 1, 0, b1 + 1)                            # :95-101
    t = dt(0)
    if b_norm:                                                               # :104-132
        t = quantile_threshold(gmag)
        if t != 0:
            if dt is f32:
                gmag = (gmag * f32(1.0 / float(f32(t / f32(n_bins))))).astype(f32)   # Mat /= s
            else:
                gmag = gmag / (t / n_bins)
    v1 = ((dt(1.0) - off1) * gmag).astype(dt)                                # :76, :144
    v2 = (off1 * gmag).astype(dt)                                            # :77, :146
    return v1, v2, b1, b2, t


def get_desc(patch, n_bins=8, init_sigma=1.4, b_norm=True, dt=f32, atan2_float=False):
    """64x64 u8 patch -> PatchTrans (4096 x n_bins), row p = column-major pixel index
    (c*64 + r) because every operand is transposed before the flatten
    (src/vgg-desc.cpp:76-77,134-148)."""
    v1, v2, b1, b2, _ = get_desc_parts(patch, n_bins, init_sigma, b_norm, dt, atan2_float)
    n = v1.size
    pt = np.zeros((n, n_bins), v1.dtype)
    # transposed flatten: p = c*rows + r
    p = np.arange(n)
    pt[p, b1.T.reshape(-1)] = v1.T.reshape(-1)
    pt[p, b2.T.reshape(-1)] = v2.T.reshape(-1)      # Bin2 assignment comes second (:143-146)
    return pt


# ---------------------------------------------------------------------------
# Distances: src/comp-fulldists.cpp:320-342 and src/comp-uprjdists.cpp:317-329
# ---------------------------------------------------------------------------
def pair_labels(indices):
    """src/comp-fulldists.cpp:247-251: Label = (3DPointID1 == 3DPointID2)."""
    indices = np.asarray(indices)
    return (indices[:, 1] == indices[:, 3]).astype(np.uint8)


def pooled_desc(filters, pt, dt=np.float64):
    """min(F * PatchTrans, 1): (nF x 4096)(4096 x 8) -> nF x 8."""
    d = filters.astype(dt) @ pt.astype(dt)
    return np.minimum(d, dt(1.0))


def full_distance(filters, pt1, pt2, dt=np.float64):
    """src/comp-fulldists.cpp:328-342 for one pair -> nF/8 ring distances."""
    l1 = pooled_desc(filters, pt2, dt) - pooled_desc(filters, pt1, dt)
    l2 = (l1 * l1).sum(axis=1)
    return l2.reshape(-1, 8).sum(axis=1)


def uprj_distance(sel_filters, pt1, pt2, dt=np.float64):
    """src/comp-uprjdists.cpp:320-329 for one pair -> (nSel*8,) = Desc1 - Desc2,
    row-major flatten [PR0 bin0..7, PR1 bin0..7, ...]."""
    d = pooled_desc(sel_filters, pt1, dt) - pooled_desc(sel_filters, pt2, dt)
    return d.reshape(-1)


def full_distances(filters, patches, indices, dt=np.float64, desc_dt=f32):
    out = np.zeros((len(indices), filters.shape[0] // 8), dt)
    cache = {}

    def desc(i):
        if i not in cache:
            cache[i] = get_desc(patches[i], dt=desc_dt)
        return cache[i]

    for k, row in enumerate(np.asarray(indices)):
        out[k] = full_distance(filters, desc(int(row[0])), desc(int(row[2])), dt)
    return out


def uprj_distances(sel_filters, patches, indices, dt=np.float64, desc_dt=f32):
    out = np.zeros((len(indices), sel_filters.shape[0] * 8), dt)
    for k, row in enumerate(np.asarray(indices)):
        out[k] = uprj_distance(sel_filters, get_desc(patches[int(row[0])], dt=desc_dt),
                               get_desc(patches[int(row[2])], dt=desc_dt), dt)
    return out


# ---------------------------------------------------------------------------
# SubtractVectorsByRows + cuda::sum: src/kernelop-opencv.cu:49-66,
# callers src/pr-learn.cpp:348-351, src/pj-learn.cpp:515-520
# ---------------------------------------------------------------------------
def hinge_pairsum(pos, neg, margin=1.0):
    """sum_i sum_j max(pos_i + margin - neg_j, 0), exact via sort + prefix sums (f64)."""
    pos = np.asarray(pos, np.float64).reshape(-1) + margin
    neg = np.sort(np.asarray(neg, np.float64).reshape(-1))
    cs = np.concatenate(([0.0], np.cumsum(neg)))
    k = np.searchsorted(neg, pos, side="left")     # neg_j < pos_i+margin
    return float(np.sum(k * pos - cs[k]))


def hinge_pairsum_bruteforce(pos, neg, margin=1.0):
    pos = np.asarray(pos, np.float64).reshape(-1, 1)
    neg = np.asarray(neg, np.float64).reshape(1, -1)
    return float(np.maximum(pos + margin - neg, 0.0).sum())


# ---------------------------------------------------------------------------
# ROC statistics: src/misc.cpp:171-264 (ComputePRStats), :266-333 (ComputePJStats)
# ---------------------------------------------------------------------------
def pr_dims(pr_params_all, w, n_channels=8):
    """nzDim / nPR / Dim bookkeeping, src/misc.cpp:182-217 (incl. the quirk
    nPR = nzDim - dup_rows/2 where dup_rows counts every row that has a twin)."""
    w = np.asarray(w).reshape(-1)
    rows = []
    for i in np.nonzero(w > 0)[0]:
        for j in range(8):
            row = pr_params_all[8 * i + j]
            if np.count_nonzero(row):
                rows.append(row)
    rows = np.array(rows, f32).reshape(-1, 3)
    dup = 0
    for i in range(len(rows)):
        same = np.all(rows == rows[i], axis=1)
        same[i] = False
        if same.any():
            dup += 1
    nz_dim = len(rows)
    n_pr = nz_dim - dup // 2
    return n_pr, n_pr * n_channels, nz_dim


def roc_stats(patch_dist, labels):
    """src/misc.cpp:227-263 == :295-332: ascending sort, float ROC counters,
    FPR at first TPR >= 0.95f, AUC = cv::contourArea of the (FPR,TPR) polygon closed
    through (1,0) (shoelace, double accumulation, |.|/2).  Ties are ordered by index
    (the reference's std::sort order is unspecified for ties)."""
    d = np.asarray(patch_dist).reshape(-1)
    lab = np.asarray(labels).reshape(-1)
    order = np.argsort(d, kind="stable")
    ls = lab[order]
    tp = np.cumsum(ls == 1).astype(f32)
    fp = np.cumsum(ls == 0).astype(f32)
    tpr = tp * f32(1.0 / float(tp[-1])) if tp[-1] > 0 else tp * f32(np.inf)
    fpr = fp * f32(1.0 / float(fp[-1])) if fp[-1] > 0 else fp * f32(np.inf)
    hit = np.nonzero(tpr >= f32(0.95))[0]
    fpr95 = f32(fpr[hit[0]]) if len(hit) else f32(-1.0)
    x = np.concatenate((fpr, [f32(1.0)])).astype(np.float64)
    y = np.concatenate((tpr, [f32(0.0)])).astype(np.float64)
    xp, yp = np.roll(x, 1), np.roll(y, 1)
    auc = abs(float(np.sum(xp * y - yp * x))) * 0.5
    return fpr95, auc


def pr_stats(pr_params_all, dists, labels, w, max_dim=-1, dt=np.float64):
    """ComputePRStats -> (nPR, Dim, nzDim, FPR95, AUC); FPR95/AUC None on MaxDim abort."""
    n_pr, dim, nz_dim = pr_dims(pr_params_all, w)
    if max_dim != -1 and dim > max_dim:
        return n_pr, dim, nz_dim, None, None
    pd = np.asarray(dists, dt) @ np.asarray(w, dt).reshape(-1)
    fpr95, auc = roc_stats(pd.astype(f32), labels)
    return n_pr, dim, nz_dim, fpr95, auc


def pj_patch_dist(W, dists, dt=np.float64):
    """colsum((W * DISTs^T)^2) over the non-zero rows of W (src/misc.cpp:269-290)."""
    W = np.asarray(W)
    nz = W[np.any(W != 0, axis=1)]
    proj = np.asarray(dists, dt) @ nz.astype(dt).T
    return nz.shape[0], (proj * proj).sum(axis=1)


def pj_stats(W, dists, labels, dt=np.float64):
    dim, pd = pj_patch_dist(W, dists, dt)
    fpr95, auc = roc_stats(pd.astype(f32), labels)
    return dim, fpr95, auc


# ---------------------------------------------------------------------------
# pr-learn: src/pr-learn.cpp:303-358
# ---------------------------------------------------------------------------
def pr_rda_steps(dists, idx_pos, idx_neg, n_pos_trn, n_neg_trn, mu, gamma, n_steps,
                 w=None, df_avg=None, t0=0, rng=None, dt=f32, samples=None):
    """Steps t0 .. t0+n_steps-1 of the l1 / non-negative RDA (single-thread order).
    Scalar factors are rounded to `dt` before use, as OpenCV's scaled convertTo does
    (SURVEY appendix B).  Returns w, df_avg, rng, min|f+1| (near-tie detector)."""
    dt = np.dtype(dt).type
    feat = dists.shape[1]
    w = np.zeros(feat, dt) if w is None else np.asarray(w, dt).copy()
    df_avg = np.zeros(feat, dt) if df_avg is None else np.asarray(df_avg, dt).copy()
    rng = CvRNG(2215) if rng is None else rng
    min_gap = np.inf
    mu = float(f32(mu))
    gamma = float(f32(gamma))
    for s in range(n_steps):
        t = t0 + s
        if samples is None:
            i_pos = rng.uniform(0, n_pos_trn)
            i_neg = rng.uniform(0, n_neg_trn)
        else:
            i_pos, i_neg = samples[s]
        diff = (dists[idx_pos[i_pos]].astype(dt) - dists[idx_neg[i_neg]].astype(dt))
        f = dt(np.dot(w.astype(np.float64), diff.astype(np.float64)))   # cv::gemm: double acc
        df_avg = df_avg * dt(t * (1.0 / (t + 1)))
        if f > -1:
            df_avg = diff * dt(1.0 / (t + 1)) + df_avg
        min_gap = min(min_gap, abs(float(f) + 1.0))
        alpha = -math.sqrt(t + 1) / gamma
        w = np.maximum(df_avg * dt(alpha) + dt(alpha * mu), dt(0))
    return w, df_avg, rng, min_gap


def pr_validate(pos_val, neg_val, w, mu, dt=np.float64):
    """src/pr-learn.cpp:340-358 -> (LossVal, Regul)."""
    w = np.asarray(w, dt).reshape(-1)
    pd = np.asarray(pos_val, dt) @ w
    nd = np.asarray(neg_val, dt) @ w
    loss = hinge_pairsum(pd, nd) / (len(pd) * len(nd))
    return loss, float(f32(mu)) * float(np.abs(w).sum())


# ---------------------------------------------------------------------------
# pj-learn: src/pj-learn.cpp:305-527
# ---------------------------------------------------------------------------
def pj_sample_batch(rng, n_pos_trn, n_neg_trn, sz_batch=200):
    """Interleaved pos,neg draws (src/pj-learn.cpp:311-316)."""
    ip = np.empty(sz_batch, np.int64)
    ineg = np.empty(sz_batch, np.int64)
    for k in range(sz_batch):
        ip[k] = rng.uniform(0, n_pos_trn)
        ineg[k] = rng.uniform(0, n_neg_trn)
    return ip, ineg


def pj_margins(W, X, dt=np.float64):
    """colsum((W X^T)^2): src/pj-learn.cpp:332-365."""
    p = np.asarray(X, dt) @ np.asarray(W, dt).T
    return (p * p).sum(axis=1)


def pj_grad(pos, neg, pos_dist, neg_dist, dt=np.float64):
    """src/pj-learn.cpp:367-418: dLoss = sum_i [ nViol_i p_i p_i^T - sum_{k viol} n_k n_k^T ]
    = X^T diag(c) X with X=[pos;neg].  -> dLoss, c_pos, c_neg."""
    viol = (np.asarray(pos_dist, f32)[:, None] + f32(1.0)) > np.asarray(neg_dist, f32)[None, :]
    c_pos = viol.sum(axis=1).astype(dt)
    c_neg = -viol.sum(axis=0).astype(dt)
    P = np.asarray(pos, dt)
    N = np.asarray(neg, dt)
    d_loss = (P.T * c_pos) @ P + (N.T * c_neg) @ N
    return d_loss, c_pos, c_neg


def pj_prox(A, dt=f32):
    """src/pj-learn.cpp:434-490: eigen-decomposition (LAPACKE_ssyevr, ascending),
    clip at 0, A = V diag(e+) V^T, W = rows sqrt(e+_k) v_k^T for e+_k != 0."""
    from scipy.linalg import lapack
    A = np.asarray(A)
    if np.dtype(dt) == np.float32:
        ev, evec, _m, _isuppz, info = lapack.ssyevr(A.astype(f32), compute_v=1, range="A", lower=0)
        assert info == 0
    else:
        ev, evec, _m, _isuppz, info = lapack.dsyevr(A.astype(np.float64), compute_v=1, range="A", lower=0)
        assert info == 0
    e = np.maximum(ev, 0)
    A_new = (evec * e) @ evec.T
    keep = e != 0
    W = (evec[:, keep] * np.sqrt(e[keep])).T
    if W.shape[0] == 0:
        W = np.zeros_like(A_new)
    return A_new.astype(dt), W.astype(dt), e


def pj_rda_steps(dists, idx_pos, idx_neg, n_pos_trn, n_neg_trn, mu, gamma, n_steps,
                 sz_batch=200, dt=f32, state=None):
    """Steps of the nuclear-norm RDA.  state = dict(W, A, dfAvg, t, rng) or None."""
    dt = np.dtype(dt).type
    D = dists.shape[1]
    if state is None:
        state = dict(W=np.zeros((D, D), dt), A=np.zeros((D, D), dt),
                     dfAvg=np.zeros((D, D), dt), t=0, rng=CvRNG(2215))
    W, A, df_avg, t, rng = state["W"], state["A"], state["dfAvg"], state["t"], state["rng"]
    mu = float(f32(mu))
    gamma = float(f32(gamma))
    for _ in range(n_steps):
        ip, ineg = pj_sample_batch(rng, n_pos_trn, n_neg_trn, sz_batch)
        pos = dists[idx_pos[ip]].astype(dt)
        neg = dists[idx_neg[ineg]].astype(dt)
        pd = pj_margins(W, pos, dt).astype(f32)
        nd = pj_margins(W, neg, dt).astype(f32)
        d_loss, _, _ = pj_grad(pos, neg, pd, nd, dt)
        beta = f32(1.0) / f32(sz_batch * sz_batch * (t + 1))            # :422 (float)
        df_avg = (df_avg * dt(t / (t + 1)) + d_loss.astype(dt) * dt(beta)).astype(dt)
        div = -math.sqrt(t + 1.0) / gamma                               # :426-428
        A = (df_avg + dt(mu) * np.eye(D, dtype=dt)) * dt(div)
        A = (dt(0.5) * (A + A.T)).astype(dt)                            # :432
        A, W, _ = pj_prox(A, dt)
        t += 1
    return dict(W=W, A=A, dfAvg=df_avg, t=t, rng=rng)


def pj_validate(pos_val, neg_val, W, A, mu, dt=np.float64):
    """src/pj-learn.cpp:501-527 -> (LossVal, Regul = mu * trace(A))."""
    pd = pj_margins(W, pos_val, dt)
    nd = pj_margins(W, neg_val, dt)
    loss = hinge_pairsum(pd, nd) / len(pd) / len(nd)
    return loss, float(f32(mu)) * float(np.trace(np.asarray(A, np.float64)))


# ---------------------------------------------------------------------------
# Synthetic Brown-shaped data; same layout as conv-impatches
# (src/conv-impatches.cpp:130-134,183-189): Patches u8 Px64x64, Indices i32 Nx4.
# ---------------------------------------------------------------------------
def synth_patches(n_points, per_point=3, seed=20240601, size=64):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float64)
    patches = np.zeros((n_points * per_point, size, size), np.uint8)
    point_id = np.repeat(np.arange(n_points), per_point).astype(np.int32)
    for pt in range(n_points):
        nb = rng.integers(6, 11)
        blobs = [(rng.uniform(4, 60), rng.uniform(4, 60), rng.uniform(2, 10), rng.uniform(2, 10),
                  rng.uniform(0, np.pi), rng.uniform(-1, 1)) for _ in range(nb)]
        for k in range(per_point):
            dx, dy = rng.uniform(-2, 2, 2)
            rot = np.deg2rad(rng.uniform(-5, 5))
            c, s = np.cos(rot), np.sin(rot)
            xr = c * (xx - 31.5) - s * (yy - 31.5) + 31.5 + dx
            yr = s * (xx - 31.5) + c * (yy - 31.5) + 31.5 + dy
            f = np.zeros((size, size))
            for (cx, cy, sx, sy, th, a) in blobs:
                ct, st = np.cos(th), np.sin(th)
                u = ct * (xr - cx) + st * (yr - cy)
                v = -st * (xr - cx) + ct * (yr - cy)
                f += a * np.exp(-0.5 * ((u / sx) ** 2 + (v / sy) ** 2))
            gain, bias = rng.uniform(0.8, 1.2), rng.uniform(-10, 10)
            img = 128 + gain * 70 * f + bias + rng.normal(0, 4, (size, size))
            patches[pt * per_point + k] = np.clip(np.rint(img), 0, 255).astype(np.uint8)
    return patches, point_id


def synth_pairs(point_id, n_pairs, seed=20240602):
    """Exactly 50 % positives (same 3-D point, different patch) / 50 % negatives."""
    rng = np.random.default_rng(seed)
    n_patch = len(point_id)
    per = int(np.sum(point_id == point_id[0]))
    idx = np.zeros((n_pairs, 4), np.int32)
    n_pos = n_pairs // 2
    for k in range(n_pairs):
        a = int(rng.integers(0, n_patch))
        if k < n_pos:
            base = (a // per) * per
            b = base + (a - base + int(rng.integers(1, per))) % per
        else:
            b = int(rng.integers(0, n_patch))
            while point_id[b] == point_id[a]:
                b = int(rng.integers(0, n_patch))
        idx[k] = (a, point_id[a], b, point_id[b])
    rng.shuffle(idx, axis=0)
    return idx
