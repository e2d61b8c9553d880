#!/usr/bin/env python
"""Golden vectors from ALL 15 data/MiddEval3/trainingH pairs (BASELINE.json configs[1]): one crop per scene,
searched with the reference's own arithmetic executed through cv2 (make_golden.ref_right, a line-by-line
transliteration of BlockSearch.cpp:88-179).  Two parameter sets per scene: main.cpp's default window
(blockSize 17, smoothFactor 0.9) and a 9x9 window without the smoothing rule.  Crops keep the run short
(the cv2 transliteration evaluates ~1e5 candidates/s).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_middeval.py
The outputs (middeval_*.npz, kind "right") are committed; the parity tests read only the .npz files."""
import os
import sys

import cv2
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import REF, OUT, ref_right, crop  # noqa: E402

SCENES = ["Adirondack", "ArtL", "Jadeplant", "Motorcycle", "MotorcycleE", "Piano", "PianoL", "Pipes", "Playroom",
          "Playtable", "PlaytableP", "Recycle", "Shelves", "Teddy", "Vintage"]


def main():
    for i, sc in enumerate(SCENES):
        L = cv2.imread(f"{REF}/data/MiddEval3/trainingH/{sc}/im0.png")
        R = cv2.imread(f"{REF}/data/MiddEval3/trainingH/{sc}/im1.png")
        assert L is not None and R is not None, sc
        h, w = R.shape[:2]
        # a different place in every scene; every third crop touches the image's right / bottom border
        y0 = (37 * i + 50) % (h - 60)
        x0 = (113 * i + 90) % (w - 140)
        if i % 3 == 0:
            x0 = w - 88
        if i % 3 == 1:
            y0 = h - 34
        ch, cwr, cwl = 34, 64, 88
        Rc = crop(R, y0, x0, ch, min(cwr, w - x0))
        Lc = crop(L, y0, x0, ch, min(cwl, L.shape[1] - x0))
        for tag, block, mx, s in (("b17_s09", 17, 24, 0.9), ("b9_s1", 9, 24, 1.0)):
            D, mb = ref_right(Lc, Rc, block, 0, mx, s)
            name = f"middeval_{sc.lower()}_{tag}"
            np.savez_compressed(os.path.join(OUT, name + ".npz"), kind="right", left=Lc, right=Rc, block=block, min_d=0,
                                max_d=mx, smooth=s, var_block=0, thres=19.0, out=D, max_block=mb)
            print(name, Lc.shape, Rc.shape, "nonzero", int((D != 0).sum()), flush=True)


if __name__ == "__main__":
    main()
