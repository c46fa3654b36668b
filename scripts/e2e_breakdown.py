import importlib, sys, time, ctypes as C
sys.path.insert(0, '.')
import torch
pkg = importlib.import_module("raytracer-challenge-rs_b200"); scenes = importlib.import_module("raytracer-challenge-rs_b200.scenes")
name = sys.argv[1] if len(sys.argv) > 1 else "reflect_refract"
W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
N = int(sys.argv[4]) if len(sys.argv) > 4 else 20
w, c = getattr(scenes, name)(pkg.scene_api(), width=W, height=H)
host = torch.empty(W*H*3, dtype=torch.float64).pin_memory()
cam = pkg.camera_to_rtc(c); opts = pkg.make_opts(); st = pkg.RtcStats(); lib = pkg.cuda_lib()
def step(T):
    t0 = time.perf_counter(); f = pkg.FlatScene(w)
    t1 = time.perf_counter(); ds = pkg.DeviceScene(f)
    t2 = time.perf_counter(); pkg._check(lib.rtc_render(ds.h, C.byref(cam), C.byref(opts), host.data_ptr(), C.byref(st)))
    t3 = time.perf_counter(); ds.close()
    t4 = time.perf_counter()
    T.append((t1-t0, t2-t1, t3-t2, t4-t3, st.kernel_ms, st.total_ms))
T = []
for _ in range(3): step(T)
T = []
for _ in range(N): step(T)
import numpy as np
a = np.array(T).mean(0)
print("flatten %.3f ms  scene_create %.3f ms  rtc_render %.3f ms (kernel %.3f, total_ms %.3f)  destroy %.3f ms" % (a[0]*1e3, a[1]*1e3, a[2]*1e3, a[4], a[5], a[3]*1e3))
# resident scene: render only
ds = pkg.DeviceScene(pkg.FlatScene(w))
for _ in range(3): lib.rtc_render(ds.h, C.byref(cam), C.byref(opts), host.data_ptr(), C.byref(st))
t0 = time.perf_counter()
for _ in range(N): lib.rtc_render(ds.h, C.byref(cam), C.byref(opts), host.data_ptr(), C.byref(st))
print("resident scene: rtc_render %.3f ms/frame (kernel %.3f)" % ((time.perf_counter()-t0)*1e3/N, st.kernel_ms))
