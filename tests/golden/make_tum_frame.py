#!/usr/bin/env python
"""Generates tests/golden/tum_frame.npz from the one real RGB-D frame the reference ships as a test fixture
(src/vslam/src/odometry/test/resource/rgb.png, depth.png: a TUM frame, 640x480, 16-bit depth with scale 1/5000,
odometry/test/test_alignment_se3.cpp:99; a third of its depth pixels are invalid).  Two overlapping 320x240 crops of
it (offset by a few pixels) make a frame pair with real texture and real depth holes; the OpenCV pyramid / blur / Sobel
outputs of the first crop pin the image operations on real imagery.  Run in the build container (needs cv2 and
/root/reference); the tests only read the .npz.

    python tests/golden/make_tum_frame.py
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RES = "/root/reference/src/vslam/src/odometry/test/resource"


def main():
    bgr = cv2.imread(os.path.join(RES, "rgb.png"), cv2.IMREAD_COLOR)
    depth = cv2.imread(os.path.join(RES, "depth.png"), cv2.IMREAD_UNCHANGED)
    assert bgr.shape == (480, 640, 3) and depth.shape == (480, 640) and depth.dtype == np.uint16
    gray = cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)  # == COLOR_RGB2GRAY of the RGB image the node receives
    W, H = 320, 240
    offs = [(80, 20), (83, 22)]  # a region where 40 % of the depth pixels are invalid
    data = {"cv2_version": np.array(cv2.__version__), "offsets": np.array(offs)}
    for k, (x0, y0) in enumerate(offs):
        data[f"gray{k}"] = np.ascontiguousarray(gray[y0:y0 + H, x0:x0 + W])
        data[f"depth{k}"] = np.ascontiguousarray(depth[y0:y0 + H, x0:x0 + W])
    img = data["gray0"]
    pyr = [img, cv2.pyrDown(img)]
    pyr.append(cv2.pyrDown(pyr[1]))
    for l, p in enumerate(pyr):
        if l:
            data[f"pyr{l}"] = p
        blur = cv2.GaussianBlur(p, (3, 3), 0, 0, borderType=cv2.BORDER_DEFAULT)
        data[f"dx{l}"] = cv2.Sobel(blur, cv2.CV_16S, 1, 0, ksize=3)
        data[f"dy{l}"] = cv2.Sobel(blur, cv2.CV_16S, 0, 1, ksize=3)
    np.savez_compressed(os.path.join(HERE, "tum_frame.npz"), **data)
    print("wrote tum_frame.npz", {k: getattr(v, "shape", None) for k, v in data.items()})


if __name__ == "__main__":
    main()
