"""Copies the reference's shipped data that the full-size parity tests need into tests/data/ (run once in the build
container, where /root/reference exists; the copies are committed because the GPU box has no reference tree):

  data/MiddEval3/trainingH/*/{im0.png, im1.png, calib.txt, mask0nocc.png, mask1nocc.png}   15 half-resolution pairs
  data/MiddEval3/trainingH/{ArtL,Teddy}/disp{0,1}GT.pfm                                   the only shipped ground truth
  results/Rectified/trainingH/*/rectified{Left,Right}.png      11 outputs of the reference's Rectification = the real
                                                              inputs of BlockSearch on main.cpp's path (2 of unequal size)
  results/PerceptualWindowSearch/test_result.png               a block-search result for Teddy (soft golden)

These are data fixtures (inputs and ground truth), not reference source code."""
import os
import shutil
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def cp(src, dst):
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    shutil.copyfile(src, dst)


n = 0
t = os.path.join(REF, "data", "MiddEval3", "trainingH")
for scene in sorted(os.listdir(t)):
    for f in sorted(os.listdir(os.path.join(t, scene))):
        cp(os.path.join(t, scene, f), os.path.join(HERE, "MiddEval3", "trainingH", scene, f)); n += 1
r = os.path.join(REF, "results", "Rectified", "trainingH")
for scene in sorted(os.listdir(r)):
    for f in ("rectifiedLeft.png", "rectifiedRight.png"):
        cp(os.path.join(r, scene, f), os.path.join(HERE, "Rectified", "trainingH", scene, f)); n += 1
cp(os.path.join(REF, "results", "PerceptualWindowSearch", "test_result.png"), os.path.join(HERE, "PerceptualWindowSearch", "test_result.png")); n += 1
print(f"{n} files copied")
