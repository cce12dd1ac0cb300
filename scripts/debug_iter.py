import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
from oracle import gn_oracle as G
g = dict(np.load('tests/golden/reference_outputs.npz'))
st = synth.make_decoder_state(0)
dev = DeviceDecoder(st, 0); fd = G.FoldedDecoder(st)
import itertools
for (name, style), prec in itertools.product((('kitti0','kitti'),('kitti1','kitti'),('redwood1','redwood')), ('fp32','tensor')):
    dev.set_precision(prec); print('=====', prec)
    p = name + '_'
    cfg = synth.optimizer_config(style, num_depth_samples=50, num_iterations=1)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg)); opt.set_debug(True)
    out = opt.reconstruct_object(g[p+'t_cam_obj'], g[p+'pts'], g[p+'rays'], g[p+'depth'], g[p+'code0'])
    it = opt.get_iteration(0)
    prm = G.Params(cfg)
    t_oc = torch.inverse(torch.from_numpy(g[p+'t_cam_obj']))
    o = G.gn_iteration(fd, prm, t_oc, torch.from_numpy(g[p+'code0']), torch.from_numpy(g[p+'pts']), torch.from_numpy(g[p+'rays']), torch.from_numpy(g[p+'depth']))
    ren = o['render']
    print(name, 'counts gpu', it['n_inball'], it['n_band'], it['n_keep'], 'oracle', ren['valid_ray'].shape[0], ren['band_ray'].shape[0], ren['keep_ray'].shape[0])
    print(' loss gpu', out.loss, 'oracle', float(o['loss']), 'golden', float(g[p+'it1_loss']))
    H, Hr = it['H'], o['H'].numpy()
    d = np.abs(H - Hr); i = np.unravel_index(d.argmax(), d.shape)
    print(' H maxdiff', d.max(), 'at', i, H[i], Hr[i], 'Hmax', np.abs(Hr).max())
    print(' H pose block diff', np.abs(H[:7,:7]-Hr[:7,:7]).max(), ' code block', np.abs(H[7:,7:]-Hr[7:,7:]).max(), 'cross', np.abs(H[:7,7:]-Hr[:7,7:]).max())
    b, br = it['b'], o['b'].numpy()
    print(' b diff', np.abs(b-br).max(), np.abs(br).max(), 'pose', np.abs(b[:7]-br[:7]).max())
    print(' b gpu', b[:7]); print(' b orc', br[:7])
    dx, dr = it['dx'], o['dx'].numpy()
    print(' dx diff', np.abs(dx-dr).max(), np.abs(dr).max()); print(' dx gpu', dx[:7]); print(' dx orc', dr[:7])
    print(' res_rot orc', float(o['res_rot']), o['J_rot'].numpy())
    # solve oracle H with gpu? check solve accuracy
    dx2 = np.linalg.solve(Hr.astype(np.float64), br.astype(np.float64))
    print(' oracle dx vs fp64 solve', np.abs(dx2-dr).max(), ' gpu dx vs fp64 solve of gpu H', np.abs(np.linalg.solve(H.astype(np.float64), b.astype(np.float64))-dx).max())
    print(' pose diff vs golden', np.abs(out.t_cam_obj-g[p+'it1_t']).max(), ' oracle vs golden', np.abs(torch.inverse(o['t_obj_cam']).numpy()-g[p+'it1_t']).max())
