#!/usr/bin/env python
"""Golden vectors for the FALS normal estimator and the loader's pixel conversions.

The reference's NormalEstimator (cpp/include/normals/NormalEstimator.h:52-176) is a sequence of OpenCV calls; OpenCV is
not vendored by the reference (FIND_PACKAGE(OpenCV), version unpinned).  This script runs the SAME call sequence with the
OpenCV that is present in this container (cv2, version recorded in the fixture) and stores inputs and outputs:

    python tests/golden/make_golden_normals.py      # writes tests/golden/normals.npz

Inputs are small so that the fixture stays a few hundred KB.  Run in the build container only (needs cv2).
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def box(src, win):
    # boxFilter(src, dst, -1, window_size_, Point(-1, -1), false): unnormalised, BORDER_DEFAULT (reflect-101)
    return cv2.boxFilter(src, -1, (win, win), anchor=(-1, -1), normalize=False)


def cache(W, H, K, win):
    """NormalEstimator::cache() (NormalEstimator.h:52-123), T = float."""
    K = K.astype(np.float64)
    fx_inv, fy_inv, cx, cy = 1.0 / K[0, 0], 1.0 / K[1, 1], K[0, 2], K[1, 2]
    row = (np.arange(W, dtype=np.float64) - cx).reshape(1, W)
    col = (np.arange(H, dtype=np.float64) - cy).reshape(H, 1)
    u = cv2.repeat(row, H, 1)
    v = cv2.repeat(col, 1, W)
    x0 = fx_inv * u
    x0_sq = cv2.multiply(x0, x0)
    y0 = fy_inv * v
    y0_sq = cv2.multiply(y0, y0)
    x0_y0 = cv2.multiply(x0, y0)
    n_sq = 1.0 + x0_sq + y0_sq
    n_sq_inv = cv2.divide(1.0, n_sq)
    x0_n = cv2.multiply(x0, n_sq_inv)
    y0_n = cv2.multiply(y0, n_sq_inv)
    M11 = box(cv2.multiply(x0_sq, n_sq_inv), win)
    M12 = box(cv2.multiply(x0_y0, n_sq_inv), win)
    M13 = box(x0_n, win)
    M22 = box(cv2.multiply(y0_sq, n_sq_inv), win)
    M23 = box(y0_n, win)
    M33 = box(n_sq_inv, win)
    det = M11 * (M22 * M33) + 2 * M12 * (M23 * M13) - (M13 * (M13 * M22) + M12 * (M12 * M33) + M23 * (M23 * M11))
    det_inv = cv2.divide(1.0, det)
    Q = [det_inv * (M22 * M33 - M23 * M23), det_inv * (M13 * M23 - M12 * M33), det_inv * (M12 * M23 - M13 * M22),
         det_inv * (M11 * M33 - M13 * M13), det_inv * (M12 * M13 - M11 * M23), det_inv * (M11 * M22 - M12 * M12)]
    f32 = lambda a: a.astype(np.float32)
    return dict(x0=f32(x0), y0=f32(y0), x0_n=f32(x0_n), y0_n=f32(y0_n), n_sq_inv=f32(n_sq_inv), Q=np.stack([f32(q) for q in Q]))


def compute(c, depth, win):
    """NormalEstimator::compute() (NormalEstimator.h:148-176), T = float."""
    tmp = cv2.divide(1.0, depth)
    z_inv = np.zeros_like(tmp)
    mask = depth != 0
    z_inv[mask] = tmp[mask]
    b1 = box(cv2.multiply(c["x0_n"], z_inv), win)
    b2 = box(cv2.multiply(c["y0_n"], z_inv), win)
    b3 = box(cv2.multiply(c["n_sq_inv"], z_inv), win)
    Q = c["Q"]
    nx = cv2.add(cv2.add(cv2.multiply(b1, Q[0]), cv2.multiply(b2, Q[1])), cv2.multiply(b3, Q[2]))
    ny = cv2.add(cv2.add(cv2.multiply(b1, Q[1]), cv2.multiply(b2, Q[3])), cv2.multiply(b3, Q[4]))
    nz = cv2.add(cv2.add(cv2.multiply(b1, Q[2]), cv2.multiply(b2, Q[4])), cv2.multiply(b3, Q[5]))
    norm = cv2.sqrt(cv2.add(cv2.add(cv2.multiply(nx, nx), cv2.multiply(ny, ny)), cv2.multiply(nz, nz)))
    with np.errstate(all="ignore"):
        return cv2.divide(nx, norm), cv2.divide(ny, norm), cv2.divide(nz, norm), b1, b2, b3


def main():
    rng = np.random.default_rng(77)
    H, W, win = 96, 128, 11
    K = np.array([[105.0, 0, 63.5], [0, 105.0, 47.5], [0, 0, 1]], np.float32)
    # a tilted plane + bumps, quantised like a TUM depth PNG (1/5000 m), with holes and a zero border strip
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float64)
    z = 1.2 + 0.002 * xx - 0.0015 * yy + 0.03 * np.sin(xx / 9.0) * np.cos(yy / 7.0)
    d16 = np.clip(np.round(z * 5000.0), 0, 65535).astype(np.uint16)
    d16[rng.random((H, W)) < 0.03] = 0
    d16[:, :7] = 0
    d16[40:60, 50:80] = 0          # a hole larger than the window: 0 / 0 normals inside
    unit = np.float32(1.0 / 5000.0)
    # ImageLoader::load_depth (ImageLoader.h:130-146): depthIn.convertTo(depth, CV_32FC1, unit_).  convertTo has no Python
    # binding; for 16U -> 32F it evaluates fma(float(src), float(alpha), 0) per pixel = one correctly rounded float product.
    depth = (d16.astype(np.float32) * unit).astype(np.float32)
    assert np.array_equal(depth, cv2.multiply(d16.astype(np.float32), np.full((H, W), unit, np.float32)))
    color8 = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    color = (color8.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)   # ImageLoader::load_color (:166-188)
    # SharpDetector.h:22-37 modifiedLaplacian on the colour image (CV_32FC3): cv::mean(FM).val[0] is the blue channel's mean
    Mk = np.array([-1, 2, -1], np.float32).reshape(3, 1)
    Gk = cv2.getGaussianKernel(3, -1, cv2.CV_32F)
    sharp = []
    for img in (color, np.ascontiguousarray(color[::-1, ::-1]), cv2.GaussianBlur(color, (5, 5), 0)):
        Lx = cv2.sepFilter2D(img, cv2.CV_32F, Mk, Gk)
        Ly = cv2.sepFilter2D(img, cv2.CV_32F, Gk, Mk)
        FM = cv2.add(np.abs(Lx), np.abs(Ly))
        sharp.append(np.float32(cv2.mean(FM)[0]))
    blurred = cv2.GaussianBlur(color, (5, 5), 0)
    c = cache(W, H, K, win)
    nx, ny, nz, b1, b2, b3 = compute(c, depth, win)
    np.savez_compressed(os.path.join(HERE, "normals.npz"), cv2_version=cv2.__version__, H=H, W=W, window=win, K=K, depth_u16=d16,
                        depth_unit=unit, color_u8=color8, depth=depth, color=color, x0=c["x0"], y0=c["y0"], x0_n=c["x0_n"],
                        y0_n=c["y0_n"], n_sq_inv=c["n_sq_inv"], Q=c["Q"], nx=nx, ny=ny, nz=nz, b1=b1, b2=b2, b3=b3, sharpness=np.array(sharp, np.float32), blurred=blurred)
    print("wrote normals.npz", cv2.__version__, "NaN normals:", int(np.isnan(nx).sum()))


if __name__ == "__main__":
    main()
