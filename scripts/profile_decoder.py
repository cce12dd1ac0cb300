"""Small driver for ncu: launches the decoder forward and forward+Jacobian kernels a few times on
row counts of a real kitti_batch32 step.  usage: python scripts/profile_decoder.py [rows_fwd rows_jac]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 479701
nj = int(sys.argv[2]) if len(sys.argv) > 2 else 79542
dev = DeviceDecoder(synth.make_decoder_state(0), 0)
g = torch.Generator(device="cuda").manual_seed(1)
def ball(n):
    x = torch.randn((n, 3), generator=g, device="cuda")
    return (x / x.norm(dim=1, keepdim=True) * torch.rand((n, 1), generator=g, device="cuda") ** (1 / 3)).contiguous()
z = torch.zeros(64, device="cuda")
xf, xj = ball(nf), ball(nj)
so, sj, jo = torch.empty(nf, device="cuda"), torch.empty(nj, device="cuda"), torch.empty((nj, 68), device="cuda")
mode_f = sys.argv[3] if len(sys.argv) > 3 else "fast"
mode_j = sys.argv[4] if len(sys.argv) > 4 else "tensor"
for i in range(3):
    dev.decode_sdf_device(z, xf, so, mode=mode_f); t1 = dev.last_kernel_ms()
    dev.decode_sdf_jacobian_device(z, xj, sj, jo, mode=mode_j); t2 = dev.last_kernel_ms()
print(f"fwd[{mode_f}] {nf} rows {t1:.3f} ms  {nf*3.67104e6/t1/1e9:.1f} TFLOP/s ; jac[{mode_j}] {nj} rows {t2:.3f} ms {nj*7.34208e6/t2/1e9:.1f} TFLOP/s")
