"""Headless image output and the metric of record.

The reference has no image writer (its result only reaches the swapchain, raytracing_renderer.odin:271-339) and
its `ssim_compare.py` needs scikit-image and a display (`cv2.imshow`, ssim_compare.py:25).  This module supplies
what a headless run needs: an 8-bit PNG writer (zlib only), a float PFM writer for the RGBA32F accumulation
image, and `ssim()` — a restatement of `skimage.metrics.structural_similarity(A, B, channel_axis=2)` with the
defaults ssim_compare.py:20 uses: 7x7 uniform window, K1 = 0.01, K2 = 0.03, data range 255 for uint8 input,
sample covariance (N / (N - 1)), mean over the interior (window radius cropped) and over the channels.
"""
from __future__ import annotations

import struct
import zlib

import numpy as np


def write_png(path: str, rgb8: np.ndarray) -> None:
    """rgb8: (H, W, 3 or 4) uint8, row 0 = top (the renderer's convention: proj[1][1] *= -1, camera.odin:80)."""
    a = np.ascontiguousarray(rgb8, np.uint8)
    h, w, c = a.shape
    assert c in (3, 4)
    raw = b"".join(b"\x00" + a[y].tobytes() for y in range(h))

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    ihdr = struct.pack(">IIBBBBB", w, h, 8, 2 if c == 3 else 6, 0, 0, 0)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", ihdr) + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def read_png(path: str) -> np.ndarray:
    """Reader for what write_png writes (8-bit RGB/RGBA, no interlace, filter 0) — enough for the SSIM CLI."""
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n", "not a PNG"
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if tag == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            assert depth == 8 and ctype in (2, 6) and interlace == 0, "unsupported PNG flavour"
            c = 3 if ctype == 2 else 4
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    out = np.zeros((h, w * c), np.uint8)
    prev = np.zeros(w * c, np.int32)
    for y in range(h):  # undo the per-row filters (0 none, 1 sub, 2 up; 3/4 only for foreign files)
        ft, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:
            cur = np.zeros(w * c, np.int32)
            for i in range(w * c):
                a_ = cur[i - c] if i >= c else 0
                b_ = prev[i]
                c_ = prev[i - c] if i >= c else 0
                if ft == 1:
                    pred = a_
                elif ft == 3:
                    pred = (a_ + b_) >> 1
                else:
                    p = a_ + b_ - c_
                    pa, pb, pc = abs(p - a_), abs(p - b_), abs(p - c_)
                    pred = a_ if (pa <= pb and pa <= pc) else (b_ if pb <= pc else c_)
                cur[i] = (line[i] + pred) & 255
        out[y] = cur
        prev = cur
    return out.reshape(h, w, c)


def write_pfm(path: str, rgb32f: np.ndarray) -> None:
    """Linear float image (H, W, 3+): PFM stores rows bottom-up, little-endian (scale -1)."""
    a = np.ascontiguousarray(rgb32f[..., :3], np.float32)
    h, w, _ = a.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(a[::-1].astype("<f4").tobytes())


def ssim(image_a: np.ndarray, image_b: np.ndarray, win_size: int = 7, data_range: float = None) -> float:
    """skimage.metrics.structural_similarity(A, B, channel_axis=2) with its defaults (ssim_compare.py:20)."""
    from scipy.ndimage import uniform_filter

    a, b = np.asarray(image_a), np.asarray(image_b)
    assert a.shape == b.shape and a.ndim == 3
    if data_range is None:
        data_range = 255.0 if a.dtype == np.uint8 else float(max(a.max(), b.max()) - min(a.min(), b.min()))
    a, b = a.astype(np.float64), b.astype(np.float64)
    k1, k2 = 0.01, 0.03
    c1, c2 = (k1 * data_range) ** 2, (k2 * data_range) ** 2
    npx = win_size ** 2
    cov_norm = npx / (npx - 1.0)  # sample covariance
    pad = (win_size - 1) // 2
    scores = []
    for ch in range(a.shape[2]):
        x, y = a[..., ch], b[..., ch]
        ux, uy = uniform_filter(x, win_size), uniform_filter(y, win_size)
        uxx, uyy, uxy = uniform_filter(x * x, win_size), uniform_filter(y * y, win_size), uniform_filter(x * y, win_size)
        vx, vy, vxy = cov_norm * (uxx - ux * ux), cov_norm * (uyy - uy * uy), cov_norm * (uxy - ux * uy)
        s = ((2 * ux * uy + c1) * (2 * vxy + c2)) / ((ux * ux + uy * uy + c1) * (vx + vy + c2))
        scores.append(s[pad:-pad, pad:-pad].mean())
    return float(np.mean(scores))
