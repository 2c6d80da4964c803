#!/usr/bin/env python
"""Generates tests/golden/cv_ops.npz: OpenCV outputs for the image operations the reference delegates to OpenCV
(core/src/Frame.cpp:264 cv::buildPyramid; :225 cv::GaussianBlur 3x3 sigma 0 BORDER_DEFAULT; :227-228 cv::Sobel
CV_16S ksize 3; src/ros/nodes/NodeRgbdAlignment.cpp:182 cv::cvtColor COLOR_RGB2GRAY).  Run in the build container (cv2 4.13 is importable there); the GPU box only reads the .npz.

    python tests/golden/make_golden.py
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def images():
    rng = np.random.default_rng(20221010)
    out = {}
    out["noise_64x48"] = rng.integers(0, 256, (48, 64), dtype=np.uint8)
    out["noise_odd_37x23"] = rng.integers(0, 256, (23, 37), dtype=np.uint8)      # odd sizes: (W+1)/2 halving
    out["saturated_32x32"] = (rng.integers(0, 2, (32, 32)) * 255).astype(np.uint8)  # extremes of every rounding
    yy, xx = np.mgrid[0:60, 0:80]
    out["smooth_80x60"] = np.clip(127 + 90 * np.sin(xx / 7.0) * np.cos(yy / 5.0) + rng.normal(0, 3, xx.shape), 0, 255).astype(np.uint8)
    out["tiny_5x4"] = rng.integers(0, 256, (4, 5), dtype=np.uint8)
    out["row_16x2"] = rng.integers(0, 256, (2, 16), dtype=np.uint8)
    return out


def main():
    data = {"cv2_version": np.array(cv2.__version__)}
    for name, img in images().items():
        data[name + "/img"] = img
        pyr = [img]
        while min(pyr[-1].shape) >= 2 and len(pyr) < 4:
            pyr.append(cv2.pyrDown(pyr[-1]))  # cv::buildPyramid == repeated pyrDown, BORDER_DEFAULT
        for l, p in enumerate(pyr[1:], 1):
            data[f"{name}/pyr{l}"] = p
        blur = cv2.GaussianBlur(img, (3, 3), 0, 0, borderType=cv2.BORDER_DEFAULT)
        data[name + "/blur"] = blur
        data[name + "/dx"] = cv2.Sobel(blur, cv2.CV_16S, 1, 0, ksize=3)
        data[name + "/dy"] = cv2.Sobel(blur, cv2.CV_16S, 0, 1, ksize=3)
    rng = np.random.default_rng(182)
    rgb = rng.integers(0, 256, (40, 56, 3), dtype=np.uint8)
    r = np.arange(0, 256, 5, dtype=np.uint8)
    grid = np.stack(np.meshgrid(r, r, r, indexing="ij"), -1).reshape(-1, 52, 3)   # every 5th value of each channel
    for name, img in (("rgb_noise_56x40", rgb), ("rgb_grid", np.ascontiguousarray(grid))):
        data[name + "/img"] = img
        data[name + "/gray_rgb"] = cv2.cvtColor(img, cv2.COLOR_RGB2GRAY)
        data[name + "/gray_bgr"] = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    np.savez_compressed(os.path.join(HERE, "cv_ops.npz"), **data)
    print("wrote", os.path.join(HERE, "cv_ops.npz"), len(data), "arrays")


if __name__ == "__main__":
    main()
