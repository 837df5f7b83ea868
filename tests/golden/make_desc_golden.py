#!/usr/bin/env python
"""cv2-driven PatchTrans fixtures: get_desc (reference src/vgg-desc.cpp:41-152) executed call for
call through the SAME OpenCV primitives the reference calls, from the `cv2` 4.13 wheel of this image.

Run HERE (needs cv2; the GPU box and the tests only read the fixture):

    python tests/golden/make_desc_golden.py

Each line of `get_desc_cv2` cites the reference line it executes.  Where the reference goes through a
cv::MatExpr that Python cannot call directly, the generator still runs OpenCV's own kernel:
  * `GAngle / AngleStep - 0.5f` (:70) is `GAngle.convertTo(dst, CV_32F, 1.0/AngleStep, -0.5)`
    (MatOp_AddEx::assign).  `cv2.normalize(NORM_MINMAX, dtype=CV_32F)` ends in exactly that
    `src.convertTo(dst, CV_32F, scale, shift)`; two sentinel elements {0, 8} pin its
    scale = float(1/AngleStep) and shift = -0.5 (see `_convert_scale`).  On this build the kernel is a
    fused multiply-add (AVX2 dispatch); the oracle restates it as one.
  * `GMag /= (T / nAngleBins)` (:131) is `GMag.convertTo(GMag, -1, 1.0/s)`: a multiplication by
    float(1.0 / double(float s)) -- the same kernel with shift 0.
  * `1.0f - Offset1.t()` (:76) is convertTo(alpha=-1, beta=1): exact either way.
atan2 (:66): the reference calls unqualified `atan2(float, float)` with only <cmath>-style headers in
scope, i.e. ::atan2(double,double) of the promoted arguments, + CV_PI in double, stored as float.  The
fixture holds that ("pt") and, because which overload a given libstdc++ exposes globally is a header
detail, the atan2f variant too ("pt_atan2f": float atan2f, promoted, + CV_PI, stored as float).
cv::magnitude (:59): with IPP (the wheel's default) it is an approximate ippsMagnitude_32f (+-2 ulp);
without IPP it is sqrt(fma(x,x,y*y)).  "pt" is generated with IPP off (what a distribution build of
OpenCV without ICV computes, bit-reproducible by the oracle), "pt_ipp" with the wheel's defaults.

Fixture `desc_kat.npz`: patches u8 [n,64,64]; for every variant v1, v2 (f32 [n,4096]) and bin1 (u8
[n,4096]) in PatchTrans row order p = x*64 + y (PatchTrans[p,bin1] = v1, PatchTrans[p,bin2] = v2, with
the Bin2 assignment second, :143-146); T (f32 [n]) the 0.8-quantile.
"""
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CV_PI = 3.1415926535897932384626433832795
f32 = np.float32


def _convert_scale(cv2, src, alpha, beta):
    """src.convertTo(dst, CV_32F, alpha, beta) through cv::normalize(NORM_MINMAX, CV_32F):
    scale = float((dmax-dmin)/(smax-smin)), shift = float(dmin) - float(smin*scale); with sentinels
    smin = 0, smax = 2^k and dmin = beta, dmax = beta + 2^k * float(alpha) both come out exact."""
    a = float(f32(alpha))
    flat = src.reshape(-1)
    top = 8.0
    while top <= float(flat.max()):
        top *= 2.0
    assert flat.min() >= 0.0
    ext = np.concatenate([flat, np.array([0.0, top], f32)]).astype(f32).reshape(1, -1)
    lo, hi = float(beta), float(beta) + top * a
    if a > 0:
        out = cv2.normalize(ext, None, alpha=lo, beta=hi, norm_type=cv2.NORM_MINMAX, dtype=cv2.CV_32F)
    else:
        raise ValueError("positive scales only")
    out = out.reshape(-1)
    assert out[-2] == f32(beta) and out.dtype == f32
    return out[:-2].reshape(src.shape).copy()


def get_desc_cv2(cv2, patch, n_bins=8, init_sigma=1.4, b_norm=True, atan2_float=False):
    """-> v1, v2 (f32 [4096]), bin1 (u8 [4096]), T; row order p = x*64 + y."""
    img = patch.astype(f32)                                                             # :44 convertTo CV_32F
    sig = float(f32(init_sigma))            # `float InitSigma` (:41) promoted to double: 1.39999997615814...
    img = cv2.GaussianBlur(img, (0, 0), sig, sigmaY=sig, borderType=cv2.BORDER_REPLICATE)   # :46
    kern = np.array([[-1, 0, 1]], f32)                                                  # :50-51
    ix = cv2.filter2D(img, cv2.CV_32F, kern, anchor=(-1, -1), delta=0,
                      borderType=cv2.BORDER_REPLICATE)                                  # :52
    iy = cv2.filter2D(img, cv2.CV_32F, kern.T.copy(), anchor=(-1, -1), delta=0,
                      borderType=cv2.BORDER_REPLICATE)                                  # :53
    gmag = cv2.magnitude(ix, iy)                                                        # :59
    gang = np.empty_like(gmag)                                                          # :64-66
    fy, fx = iy.reshape(-1), ix.reshape(-1)
    ga = gang.reshape(-1)
    if atan2_float:
        at = np.arctan2(fy, fx)                 # libm atan2f
        assert at.dtype == f32
        for i in range(ga.size):
            ga[i] = f32(float(at[i]) + CV_PI)
    else:
        for i in range(ga.size):
            ga[i] = f32(math.atan2(float(fy[i]), float(fx[i])) + CV_PI)
    step = f32(2.0 * CV_PI / float(f32(n_bins)))                                        # :69 float AngleStep
    ratio = _convert_scale(cv2, gang, 1.0 / float(step), -0.5)                          # :70
    off1 = (ratio - np.floor(ratio)).astype(f32)                                        # :72-74
    c = np.ceil(ratio - f32(1.0))                                                       # :84-91
    bin1 = np.where(c == -1.0, n_bins - 1, c).astype(np.uint8)
    T = f32(0)
    if b_norm:                                                                          # :104-132
        srt = cv2.sort(gmag.reshape(1, -1), cv2.SORT_ASCENDING).reshape(-1)             # :111-112
        n = srt.size
        aleph = f32(f32(n) * f32(0.8)) + f32(0.5)                                       # :119
        k = int(math.floor(float(aleph)))                                               # :120 cvFloor
        k = min(k, n - 1)
        k = max(k, 1)
        gamma = f32(aleph - f32(k))
        gamma = min(max(gamma, f32(0)), f32(1))
        T = f32(f32(f32(1) - gamma) * srt[k - 1]) + f32(gamma * srt[k])                 # :127-128
        if T != 0:
            s = f32(T / f32(n_bins))                                                    # :131 float / int
            gmag = _convert_scale(cv2, gmag, 1.0 / float(s), 0.0)
    w1 = (f32(1.0) - off1.T).astype(f32)                                                # :76
    w2 = off1.T.copy()                                                                  # :77
    g_t = gmag.T.copy()                                                                 # :136
    v1 = (w1 * g_t).astype(f32).reshape(-1)                                             # :144
    v2 = (w2 * g_t).astype(f32).reshape(-1)                                             # :146
    return v1, v2, bin1.T.copy().reshape(-1), T                                         # :134


def special_patches():
    """Edge cases: flat, axis/diagonal ramps (gradient angles that sit exactly on atan2's special
    values: +-0, pi/4 multiples), saturated blocks, checkerboard, single bright pixel, noise."""
    out = []
    yy, xx = np.mgrid[0:64, 0:64]
    out.append(np.full((64, 64), 128))                       # flat: T = 0, no normalisation (:131)
    out.append(np.zeros((64, 64)))
    out.append(np.full((64, 64), 255))
    out.append(xx * 4)                                       # Iy = +0 exactly, Ix > 0
    out.append(255 - xx * 4)                                 # Ix < 0, Iy = 0: atan2 = pi -> GAngle = 2 pi
    out.append(yy * 4)
    out.append(255 - yy * 4)
    out.append((xx + yy) * 2)                                # Ix == Iy: atan2 = pi/4 exactly
    out.append((xx - yy) * 2 + 126)
    out.append(((xx // 8 + yy // 8) % 2) * 255)              # checkerboard, saturated
    out.append(((xx // 16) % 2) * 255)
    z = np.zeros((64, 64)); z[31, 33] = 255; out.append(z)   # one pixel: most magnitudes ~0, many ties
    z = np.zeros((64, 64)); z[:, 32:] = 255; out.append(z)   # step edge
    rng = np.random.default_rng(777)
    out.append(rng.integers(0, 256, (64, 64)))
    out.append(rng.integers(0, 2, (64, 64)) * 255)
    out.append(np.clip(rng.normal(128, 3, (64, 64)), 0, 255))   # weak gradients: angle conditioning
    return np.array(out).astype(np.uint8)


def main():
    import cv2
    from oracle import dlco_oracle as O
    patches, _ = O.synth_patches(16, per_point=3, seed=20240601)
    patches = np.concatenate([patches[:48], special_patches()]).astype(np.uint8)
    n = len(patches)
    res = {}
    for name, ipp, at_f in (("pt", False, False), ("pt_atan2f", False, True), ("pt_ipp", True, False)):
        cv2.ipp.setUseIPP(ipp)
        v1 = np.zeros((n, 4096), f32)
        v2 = np.zeros((n, 4096), f32)
        b1 = np.zeros((n, 4096), np.uint8)
        T = np.zeros(n, f32)
        for i in range(n):
            v1[i], v2[i], b1[i], T[i] = get_desc_cv2(cv2, patches[i], atan2_float=at_f)
        res[name + "_v1"], res[name + "_v2"], res[name + "_bin1"], res[name + "_T"] = v1, v2, b1, T
    cv2.ipp.setUseIPP(True)
    # the IPP and atan2f variants differ from "pt" in few entries: keep them as sparse deltas
    out = {"patches": patches, "cv2_version": np.array(cv2.__version__),
           "pt_v1": res["pt_v1"], "pt_v2": res["pt_v2"], "pt_bin1": res["pt_bin1"], "pt_T": res["pt_T"]}
    for name in ("pt_atan2f", "pt_ipp"):
        for fld in ("v1", "v2", "bin1"):
            a, b = res["pt_" + fld].reshape(-1), res[name + "_" + fld].reshape(-1)
            idx = np.nonzero(a.view(np.uint32 if a.dtype == f32 else np.uint8) !=
                             b.view(np.uint32 if b.dtype == f32 else np.uint8))[0]
            out[name + "_" + fld + "_idx"] = idx.astype(np.int32)
            out[name + "_" + fld + "_val"] = b[idx]
        out[name + "_T"] = res[name + "_T"]
    np.savez_compressed(os.path.join(HERE, "desc_kat.npz"), **out)
    print("desc_kat.npz: %d patches, %.2f MB" % (n, os.path.getsize(os.path.join(HERE, "desc_kat.npz")) / 1e6))
    for name in ("pt_atan2f", "pt_ipp"):
        print(name, {f: len(out[name + "_" + f + "_idx"]) for f in ("v1", "v2", "bin1")})


if __name__ == "__main__":
    main()
