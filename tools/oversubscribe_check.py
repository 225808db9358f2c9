import sys, threading, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry
cam = synth.scaled_camera(synth.CAM_VGA, 2)
model, live, T = synth.frame_pair(cam, seed=3)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
D = dict(v=dev(model["vmap4"]), n=dev(model["nmap4"]), mi=dev(model["rgba"]), d=dev(live["depth"]), li=dev(live["rgba"]))
pose = np.eye(4, dtype=np.float32)
res = {}
def work(k, grid, threads, n):
    h = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    if threads: h.set_block(threads)
    h.set_grid(grid)
    s = torch.cuda.Stream()
    h.initFirstRGB(D["mi"], stream=s)
    for i in range(n):
        h.initICPModel(D["v"], D["n"], 25.0, pose, stream=s); h.initRGBModel(D["mi"], stream=s); h.initICP(D["d"], 25.0, stream=s); h.initRGB(D["li"], stream=s)
        t, R = h.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True, stream=s)
    res[k] = t
for name, cfgs in (("3 x full grid (1024 threads)", [(0, 0)] * 3), ("5 x 100 blocks of 512", [(100, 512)] * 5), ("mixed", [(0, 0), (148, 512), (60, 256), (0, 512)])):
    res.clear()
    t0 = time.time()
    ths = [threading.Thread(target=work, args=(k, g, th, 30)) for k, (g, th) in enumerate(cfgs)]
    [t.start() for t in ths]; [t.join() for t in ths]
    torch.cuda.synchronize()
    ref = res[0]
    assert all(np.array_equal(ref, r) for r in res.values())
    print(name, "ok", round(time.time() - t0, 2), "s")
