#!/usr/bin/env python
"""Golden vectors for N2 (imread / imwrite of JPEG): small files written by cv2 (the reference's codec:
libjpeg-turbo inside OpenCV) in the variants the decoder covers, plus cv2.imdecode's pixels.
Run here (cv2 present); the outputs under tests/golden/jpeg/ are committed.

    python tools/make_jpeg_goldens.py
"""
import os
import cv2
import numpy as np

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "jpeg")
VARIANTS = {
    "420_q100": [cv2.IMWRITE_JPEG_QUALITY, 100],
    "420_q75": [cv2.IMWRITE_JPEG_QUALITY, 75],
    "444_q90": [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444],
    "422_q90": [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422],
    "420_q90_optimised": [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_OPTIMIZE, 1],
    "420_q90_restart5": [cv2.IMWRITE_JPEG_QUALITY, 90, cv2.IMWRITE_JPEG_RST_INTERVAL, 5],
}


def card(h, w, seed):
    rng = np.random.default_rng(seed)
    img = np.full((h, w, 3), 255, np.uint8)
    for _ in range(40):                       # bubbles, bars and text-like strokes of an answer card
        y, x = int(rng.integers(0, h - 8)), int(rng.integers(0, w - 8))
        img[y:y + int(rng.integers(2, 8)), x:x + int(rng.integers(2, 30))] = rng.integers(0, 120, 3)
    img = cv2.GaussianBlur(img, (3, 3), 0.8)
    noise = rng.normal(0, 3, img.shape)
    return np.clip(img.astype(np.float64) + noise, 0, 255).astype(np.uint8)


def main():
    os.makedirs(OUT, exist_ok=True)
    expected = {}
    for (h, w) in [(97, 131), (64, 48), (33, 250)]:
        img = card(h, w, h * 1000 + w)
        for name, params in VARIANTS.items():
            key = f"{w}x{h}_{name}"
            enc = cv2.imencode(".jpg", img, params)[1]
            open(os.path.join(OUT, key + ".jpg"), "wb").write(enc.tobytes())
            expected[key] = cv2.imdecode(enc, cv2.IMREAD_COLOR)
        key = f"{w}x{h}_grey_q90"
        enc = cv2.imencode(".jpg", cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), [cv2.IMWRITE_JPEG_QUALITY, 90])[1]
        open(os.path.join(OUT, key + ".jpg"), "wb").write(enc.tobytes())
        expected[key] = cv2.imdecode(enc, cv2.IMREAD_COLOR)
    np.savez_compressed(os.path.join(OUT, "expected_pixels.npz"), **expected)
    print(len(expected), "files,", sum(os.path.getsize(os.path.join(OUT, f)) for f in os.listdir(OUT)), "bytes")


if __name__ == "__main__":
    main()
