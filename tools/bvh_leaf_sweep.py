"""C5 (1e6 triangles) by BVH leaf size (B2vpSceneDesc::options bits 8..10: triangles per leaf).  One GPU."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
pkg = ge.build()
import scenes
abi = pkg.abi
math = sys.argv[1] if len(sys.argv) > 1 else "fast"
for leaf in (1, 2, 3, 4, 6):
    sc = pkg.load_dict(scenes.c5(spp=256, n_leaves=500000, n_sensors=4096))
    sc.desc.options = leaf << abi.OPT_BVH_LEAF_SHIFT
    with pkg.DeviceScene(sc) as ds:
        ds.render(math=math)
        ms = min(ds.render(math=math).totals.gpu_ms for _ in range(3))
        r = ds.render(math=math)
        i = ds.info()
    print(json.dumps({"leaf": leaf, "math": math, "mpaths_s": r.totals.n_paths / ms / 1e3, "kernel_ms": ms, "bvh_nodes": i.bvh_nodes,
                      "depth": i.bvh_max_depth, "nodes_per_ray": r.totals.n_bvh_nodes / r.totals.n_intersect,
                      "tris_per_ray": r.totals.n_bvh_tris / r.totals.n_intersect}), flush=True)
