"""cost_map.py [workload] [W] [H] — per-pixel search cost (BVH nodes + prim tests + reference-box validations) of the accelerated
renderer from the kernel source compiled for the host with -DRTK_STATS: where a frame's heavy warps and CTAs are."""
import importlib, os, sys, ctypes as C, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("raytracer-challenge-rs_b200"); scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
import bench
wl = sys.argv[1] if len(sys.argv)>1 else "teapot_standin_1080p"
fn, kw, desc = bench.WORKLOADS[wl]
kw = dict(kw); kw["width"]=int(sys.argv[2]) if len(sys.argv)>2 else 480; kw["height"]=int(sys.argv[3]) if len(sys.argv)>3 else 270
world, cam = getattr(scenes, fn)(pkg.scene_api(), **kw)
flat = pkg.FlatScene(world)
lib = C.CDLL(os.environ.get('EMUL_STATS_LIB', '/tmp/libemul_stats.so'))  # g++ -O2 -std=c++17 -ffp-contract=off -fPIC -pthread -shared -DRTK_STATS -o /tmp/libemul_stats.so tests/cpu_emul/emul.cpp
camr = pkg.camera_to_rtc(cam); o = pkg.make_opts()
cost = np.zeros((cam.vsize, cam.hsize, 4), dtype=np.uint64)
rc = lib.emul_cost_map(C.byref(flat.desc), C.byref(camr), C.byref(o), cost.ctypes.data_as(C.c_void_p))
print("rc", rc)
tot = cost[...,1]+cost[...,2]+cost[...,3]
print("mean cost/pixel %.0f  median %.0f  p99 %.0f  max %d" % (tot.mean(), np.median(tot), np.percentile(tot,99), tot.max()))
idx = np.dstack(np.unravel_index(np.argsort(tot.ravel())[::-1][:10], tot.shape))[0]
for y,x in idx: print("px", x, y, "walks,bvh,prims,chain =", cost[y,x].tolist())
# 8x4 patch max vs sum
H,W = tot.shape; p = tot[:H//4*4,:W//8*8].reshape(H//4,4,W//8,8)
pm = p.max(axis=(1,3)); print("patch-max (warp cost) mean %.0f max %d ; sum of patch-max %.3g vs sum of pixels/32 %.3g" % (pm.mean(), pm.max(), pm.sum(), tot.sum()/32))
# CTA tiles 16x8
p = tot[:H//8*8,:W//16*16].reshape(H//8,8,W//16,16)
# warp cost model: per warp (8x4 patch) = patch-max; CTA = max of its 4 warps
w4 = tot[:H//8*8,:W//16*16].reshape(H//8,2,4,W//16,2,8).max(axis=(2,5))  # [ty, wy, tx, wx]
cta = w4.max(axis=(1,3))
print("CTA critical (max warp) mean %.0f p99 %.0f max %d" % (cta.mean(), np.percentile(cta,99), cta.max()))
rows = cta.max(axis=1)
print("per tile-row max:", rows.tolist()[:80])
