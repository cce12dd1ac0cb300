"""GPU box, experimental build: what the L2 -> shared-memory weight stream costs k_decoder_p3<1>.  QSP_TC_DEBUG=1 makes the
producer warps skip the bulk copies after a CTA pair's first tile (the ring stages keep the weights they hold; results
are garbage, timing only).  Same rows, same launch, best of 10; the board is power-capped, so the difference is the
energy of the stream as much as its bandwidth."""
import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import torch
    from qsp_slam_b200 import synth, build, _native as N
    with N.use_library(build.LIB_EXP):
        from qsp_slam_b200.decoder import DeviceDecoder
        dev = DeviceDecoder(synth.make_decoder_state(0), 0, calibrate=False)
        g = torch.Generator(device="cuda").manual_seed(1)
        n = 84672
        x = torch.randn((n, 3), generator=g, device="cuda")
        x = (x / x.norm(dim=1, keepdim=True) * torch.rand((n, 1), generator=g, device="cuda") ** (1 / 3)).contiguous()
        z = torch.zeros(64, device="cuda")
        s, j = torch.empty(n, device="cuda"), torch.empty((n, 68), device="cuda")
        best = 1e9
        for i in range(14):
            dev.decode_sdf_jacobian_device(z, x, s, j, mode="tensor")
            if i >= 4:
                best = min(best, dev.last_kernel_ms())
        print(f"QSP_TC_DEBUG={os.environ.get('QSP_TC_DEBUG', '0')}: k_decoder_p3<1> {n} rows {best:.4f} ms")
        dev.close()
else:
    for dbg in ("0", "1", "0", "1"):
        env = dict(os.environ, QSP_TC_DEBUG=dbg)
        print(subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True).stdout.strip(), flush=True)
