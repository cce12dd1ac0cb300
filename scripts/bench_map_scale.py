"""BASELINE config 5: a map-scale batch of 1024 kitti_batch32-shaped objects, STRONG scaling over the GPUs of one
box (objects sharded in contiguous blocks, one all-gather of the 82-float records at the end).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
      scripts/bench_map_scale.py [--objects 1024] [--batch 128]
Prints one JSON line on rank 0 (device-timed, max over ranks)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from qsp_slam_b200 import synth, sharding
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

ap = argparse.ArgumentParser()
ap.add_argument("--objects", type=int, default=1024)
ap.add_argument("--batch", type=int, default=128, help="objects per qsp_reconstruct_batch call")
ap.add_argument("--distinct", type=int, default=64, help="distinct synthetic objects (cycled)")
args = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, local)
cfg, base = synth.make_workload("kitti_batch32", state, objects=args.distinct, decode=lambda c, p: dev.decode_sdf(c, p))
objs = [base[i % args.distinct] for i in range(args.objects)]
lo, hi = sharding.partition(args.objects, world)[rank]
mine = objs[lo:hi]
opt = Optimizer(dev, ForceKeyErrorDict(cfg))

def run():
    recs = []
    for b0 in range(0, len(mine), args.batch):
        recs.append(opt.reconstruct_objects_records(mine[b0:b0 + args.batch]))
    return torch.from_numpy(np.concatenate(recs, 0)).cuda()

run()                                                     # warm-up (allocations)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
rec = run()
counts = [h - l for l, h in sharding.partition(args.objects, world)]
allrec = sharding.gather_records(rec, counts) if world > 1 else rec
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    good = int((allrec[:, 80] != 0).sum().item())
    print(json.dumps({"config": "map_scale_batch", "objects": args.objects, "n_gpus": world, "batch": args.batch,
                      "seconds": ms.item() / 1e3, "objects_per_s": args.objects / (ms.item() / 1e3), "good_objects": good,
                      "scaling": "strong", "api": "Optimizer.reconstruct_objects (host buffers) + all_gather_into_tensor of records"}))
if world > 1:
    dist.destroy_process_group()
