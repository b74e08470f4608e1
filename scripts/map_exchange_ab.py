"""A/B of the two exchanges of the sharded C4 mapping iteration (torchrun, one rank per GPU): alternating legs of
bench.mapping_leg so that drift and rank skew show up as spread, not as a difference."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import bench
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
res = {"nccl": [], "peer": [], "auto": []}
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    for ex in ("nccl", "peer", "auto"):
        res[ex].append(round(bench.mapping_leg(dev, rank, world, iters=40, warmup=5, exchange=ex)["iters_per_s"], 1))
if rank == 0:
    print(json.dumps({"world": world, "mapping_C4_iters_per_s": res}))
dist.destroy_process_group()
