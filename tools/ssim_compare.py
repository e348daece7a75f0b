#!/usr/bin/env python
"""Headless restatement of the reference's ssim_compare.py (same score, no scikit-image, no cv2.imshow).

    python tools/ssim_compare.py imageA.png imageB.png      ->  SSIM: 0.xxxxxx
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracer_b200.image_io import read_png, ssim  # noqa: E402


def main():
    ap = argparse.ArgumentParser(description="Compare two images using SSIM.")
    ap.add_argument("imageA", help="First image path (reference)")
    ap.add_argument("imageB", help="Second image path (test)")
    a = ap.parse_args()
    A, B = read_png(a.imageA)[..., :3], read_png(a.imageB)[..., :3]
    print(f"SSIM: {ssim(A, B):.6f}")


if __name__ == "__main__":
    main()
