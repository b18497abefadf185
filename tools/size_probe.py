import ctypes, os, sys
ROOT = os.getcwd(); sys.path.insert(0, ROOT)
import rtb200_loader
pkg = rtb200_loader.load()
rt = ctypes.CDLL("libcudart.so")
dev = ctypes.c_void_p(); assert rt.cudaMalloc(ctypes.byref(dev), ctypes.c_size_t(4 * 3840 * 2160)) == 0
for (W, H) in ((3840, 2160), (2716, 1528), (1920, 1080), (3840, 1080)):
    w = pkg.World(os.path.join(ROOT, "scenes", "world16.json"), width=W, height=H, depth=9)
    s = pkg.Scene(w)
    for env in ("1", "0"):
        os.environ["RT_BEAM_SHADOW"] = env
        ms = []
        for i in range(12):
            _, st = s.render(flags=pkg.RT_RENDER_REBUILD_BVH | pkg.RT_RENDER_CANVAS_DEVICE, out_ptr=dev.value)
            ms.append(st)
        f = lambda k: sorted(x[k] for x in ms[6:])[3]
        st = ms[-1]
        print("%dx%d beam=%s: hits %d | k_tile<1> %.3f k_shadow %.3f -> per Mhit: tile %.4f shadow %.4f" % (W, H, env, st["n_hits"], f("ms_primary"), f("ms_shadow"), f("ms_primary") / (W * H / 1e6), f("ms_shadow") / (st["n_hits"] / 1e6)))
    s.close()
