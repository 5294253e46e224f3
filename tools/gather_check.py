"""Multi-GPU check of ray-sharded inference + the final gather (SURVEY.md 8e "Inference", eval.py:186-205 sharded):
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 tools/gather_check.py
Every rank renders its contiguous pixel block of a 200x200 view (render_image -> shard_range, gather_rays); the
gathered image must equal, bit for bit, the image rank 0 renders alone (rays are independent; every point's arithmetic is
the same whatever tile it lands in)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
from nerf_fl_b200 import synthetic as syn, parallel
from nerf_fl_b200.nerf import NeRF, PosEmbedding
from nerf_fl_b200.inference import render_image

models = {"coarse": NeRF("coarse"), "fine": NeRF("fine", encode_appearance=True, encode_transient=True, beta_min=0.03)}
models["coarse"].load_state_dict(syn.make_field_params(41, "coarse"))
models["fine"].load_state_dict(syn.make_field_params(42, "fine", True, True))
models = {k: m.to(dev) for k, m in models.items()}
emb = {"xyz": PosEmbedding(9, 10), "dir": PosEmbedding(3, 4),
       "a": torch.nn.Embedding.from_pretrained(syn.make_embedding(31, 100, 48)).to(dev),
       "t": torch.nn.Embedding.from_pretrained(syn.make_embedding(32, 100, 16)).to(dev)}
H = W = 200
Kc, c2w = syn.view_camera(H, W, view=3)
sharded = render_image(models, emb, H, W, Kc, c2w, 7, 64, 64, chunk=4096, white_back=True)
# the same image on one rank: make dist_info() report a single rank for the duration of the call
real = parallel.dist_info
import nerf_fl_b200.inference as inf
inf.dist_info = lambda: (0, 1)
inf.gather_rays = lambda local_, n: local_
alone = render_image(models, emb, H, W, Kc, c2w, 7, 64, 64, chunk=4096, white_back=True)
ok = all(torch.equal(sharded[k], alone[k]) for k in ("rgb", "depth", "img_u8"))
lo, hi = parallel.shard_range(H * W, rank, world)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"world {world}: {H}x{W} view, rank 0 rendered pixels [{lo},{hi}); gathered image == single-rank image bit for bit on every rank: {bool(flag.item())}; "
          f"rgb mean {sharded['rgb'].mean().item():.6f}", flush=True)
    print("GATHER CHECK", "PASSED" if flag.item() else "FAILED", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
