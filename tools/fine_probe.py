import os, sys
sys.path.insert(0, os.getcwd())
os.environ["RT_BUILD_TIMERS"] = "1"
import rtb200_loader
pkg = rtb200_loader.load()
w = pkg.World("scenes/world16.json"); s = pkg.Scene(w)
for _ in range(10): s.build_bvh()
st = [s.debug_get(pkg.RT_DEBUG_BUILD_LEVEL_NS + 16 + i) for i in range(8)]
names = ["children (phase A)", "warp reduce", "syncthreads+btot", "cluster sync", "block prefix", "scan+syncthreads", "write nodes", "cluster sync"]
print("level 2 of world16, us:", ", ".join("%s %.2f" % (n, (st[i + 1] - st[i]) / 1e3) for i, n in enumerate(names[:7])))
