import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
dev = DeviceDecoder(synth.make_decoder_state(0), 0)
rng = np.random.default_rng(0)
worst = 0
for t in range(8):
    x = synth._ball_points(rng, 250000); z = (rng.standard_normal(64) * (0.0 if t == 0 else 0.1)).astype(np.float32)
    a = dev.decode_sdf(z, x, mode="fast"); b = dev.decode_sdf(z, x, mode="tensor")
    e = np.abs(a - b); near = np.abs(b) < 0.02
    print(f"trial {t}: max err {e.max():.2e} rms {np.sqrt((e**2).mean()):.2e} max err where |s|<0.02: {e[near].max():.2e} p99.99 {np.quantile(e, 0.9999):.2e}")
    worst = max(worst, e.max())
print("worst", worst)
