import sys, os, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import dustraytracer_b200 as drt
from run_configs import soup
n = int(sys.argv[1])
s = drt.Scene.from_triangles(soup(n))
t0 = time.perf_counter(); drt.BVHBuilder().buildIterative(s); t1 = time.perf_counter()
print(s.stats()["build_ms"], t1 - t0, s.stats()["wide_nodes"], s.stats()["bvh_levels"])
