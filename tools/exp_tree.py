"""CPU experiment: tree quality of the own SBVH builder measured as node-pair / triangle visits per ray of the
path-traced workload (instrumented oracle, same camera/scene/depth as bench.py). No GPU needed.
usage: python tools/exp_tree.py [atrium|bathroom] [ref]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench, oracle
from gpupathtracer_b200 import api, make_rng_table

name = sys.argv[1] if len(sys.argv) > 1 else "atrium"
wl = bench.WORKLOADS[name]
scene = bench.build_scene(wl)
t0 = time.time()
if "ref" in sys.argv:
    nodes, refs = oracle.ref_build_sbvh(scene.tris_abs)
    st = {}
else:
    nodes, refs, st = api.build_bvh_host(scene.tris)
dt = time.time() - t0
scene.set_bvh(nodes, refs)
w, h = 240, 136
_, _, _, tot = bench.cpu_reference_run(scene, 2, 0, w, h, wl)
nc, ns = tot["closest_rays"], tot["shadow_rays"]
print(f"{name} build {dt:.2f}s nodes {len(nodes)} refs {len(refs)} sah {st.get('sah_cost')}  closest: pairs/ray {(tot['node_pairs']-tot['shadow_node_pairs'])/nc:.3f} "
      f"tris/ray {(tot['tri_tests']-tot['shadow_tri_tests'])/nc:.3f}  shadow: pairs/ray {tot['shadow_node_pairs']/ns:.3f} tris/ray {tot['shadow_tri_tests']/ns:.3f}")
