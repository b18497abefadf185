"""Times the parts of a multi-GPU step with CUDA events: frame, all-gather, de-interleave."""
import os, sys, torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
w = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=3840, height=2160, depth=9)
s = pkg.Scene(w, device=lr)
s.set_stream(torch.cuda.current_stream().cuda_stream)
fl = pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE | pkg.RT_RENDER_TILED_OUTPUT | pkg.RT_RENDER_ASYNC
p = s.params(flags=fl, rank=rank, world=world, tile_w=32, tile_h=16)
slab = s.tiled_slab_pixels(p)
slab_t = torch.zeros(slab, dtype=torch.int32, device="cuda")
gathered = torch.zeros(world * slab, dtype=torch.int32, device="cuda")
canvas = torch.zeros(3840 * 2160, dtype=torch.int32, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
acc = [0.0, 0.0, 0.0]
for it in range(13):
    dist.barrier(); torch.cuda.synchronize()
    ev[0].record()
    s.render(flags=fl, rank=rank, world=world, tile_w=32, tile_h=16, out_ptr=slab_t.data_ptr())
    ev[1].record()
    dist.all_gather_into_tensor(gathered, slab_t)
    ev[2].record()
    s.untile(p, gathered.data_ptr(), canvas.data_ptr())
    ev[3].record()
    torch.cuda.synchronize(); s.finish()
    if it >= 3:
        for k in range(3): acc[k] += ev[k].elapsed_time(ev[k + 1]) / 10
if rank == 0:
    print("N=%d frame %.3f ms, all_gather %.3f ms, untile %.3f ms" % (world, *acc))
dist.barrier(); s.close()
