"""GPU check of the tensor-core decoder kernels against the oracle (fp32 and fp64)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from oracle import gn_oracle as G
st = synth.make_decoder_state(0)
dev = DeviceDecoder(st, 0)
fd, fd64 = G.FoldedDecoder(st), G.FoldedDecoder(st, torch.float64)
rng = np.random.default_rng(3)
for n in (1, 64, 65, 200, 5000):
    x = synth._ball_points(rng, n); z = (rng.standard_normal(64) * 0.1).astype(np.float32)
    ref64 = G.decode_sdf(fd64, torch.from_numpy(z), torch.from_numpy(x)).numpy()
    ref32 = G.decode_sdf(fd, torch.from_numpy(z), torch.from_numpy(x)).numpy()
    line = f"n={n:6d} fp32-oracle err {np.abs(ref32-ref64).max():.2e}"
    for mode in ("fp32", "tensor", "fast"):
        got = dev.decode_sdf(z, x, mode=mode)
        line += f" | {mode} {np.abs(got-ref64).max():.2e}"
    print(line, flush=True)
    if n <= 5000:
        s64, g64 = G.decoder_forward_jacobian(fd64, torch.from_numpy(z), torch.from_numpy(x))
        for mode in ("fp32", "tensor"):
            s, jac = dev.decode_sdf_jacobian(z, x, mode=mode)
            d = np.abs(jac - g64.numpy()).max(1)
            print(f"    jac {mode}: sdf err {np.abs(s-s64.numpy()).max():.2e} row err median {np.median(d):.2e} max {d.max():.2e} bad rows(>1e-4) {(d>1e-4).sum()}", flush=True)
# timing
for n in (480000, 80000):
    x = torch.from_numpy(synth._ball_points(rng, n)).cuda(); zc = torch.zeros(64, device="cuda")
    so = torch.empty(n, device="cuda"); jo = torch.empty((n, 68), device="cuda")
    for mode in ("fast", "tensor", "fp32"):
        for _ in range(3): dev.decode_sdf_device(zc, x, so, mode=mode)
        t = dev.last_kernel_ms(); print(f"fwd {mode} n={n}: {t:.3f} ms {n*3.67104e6/t/1e9:.1f} TFLOP/s-equivalent", flush=True)
    for mode in ("tensor", "fp32"):
        for _ in range(3): dev.decode_sdf_jacobian_device(zc, x, so, jo, mode=mode)
        t = dev.last_kernel_ms(); print(f"jac {mode} n={n}: {t:.3f} ms {n*7.34208e6/t/1e9:.1f} TFLOP/s-equivalent", flush=True)
