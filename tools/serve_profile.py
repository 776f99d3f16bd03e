"""cProfile of the single-scan serving path (host side); run on the GPU box."""
import cProfile, os, pstats, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
from ptv3_b200 import infer, synth
dev = torch.device("cuda", 0)
model, _ = bench.build_model(torch.bfloat16, dev)
coord, feat = synth.pig_scan(100000, scale=4.1, seed=3)
c, f = torch.from_numpy(coord).to(dev), torch.from_numpy(feat).to(dev)
grid = 0.02 / float(np.linalg.norm(coord - coord.mean(0), axis=1).max())
for _ in range(5):
    infer.detect_keypoints(model, c, f, grid).cpu()
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    infer.detect_keypoints(model, c, f, grid).cpu()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
