"""OpenCV's SIMD chain (NV12 -> RGB, bilinear resize, float normalise) on N host threads; prints one JSON line.
Child of bench.py's cpu_baseline leg: a stand-in for the AVX2 rows of the real libyuv, NOT a parity arm (OpenCV's
fixed point differs from libyuv's in the last bits). Usage: simd_proxy.py W H DW DH THREADS SECONDS"""
import json
import sys
import threading
import time

import numpy as np


def main():
    w, h, dw, dh, threads = (int(x) for x in sys.argv[1:6])
    seconds = float(sys.argv[6])
    import cv2
    cv2.setNumThreads(1)
    rng = np.random.default_rng(7)
    frame = rng.integers(0, 256, size=(h * 3 // 2, w), dtype=np.uint8)
    mean = np.float32(127.5)
    std = np.float32(127.5)
    counts = [0] * threads

    def work(i):
        t_end = time.perf_counter() + seconds
        while time.perf_counter() < t_end:
            rgb = cv2.cvtColor(frame, cv2.COLOR_YUV2RGB_NV12)
            small = cv2.resize(rgb, (dw, dh), interpolation=cv2.INTER_LINEAR)
            _ = (small.astype(np.float32) - mean) / std
            counts[i] += 1

    ths = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    print(json.dumps({"frames_per_s": sum(counts) / dt,
                      "what": "OpenCV %s cvtColor(YUV2RGB_NV12) + resize(INTER_LINEAR) + float normalise, %d threads, %.0f s"
                              % (cv2.__version__, threads, seconds)}))


if __name__ == "__main__":
    main()
