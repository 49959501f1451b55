import os, sys, time, ctypes as C
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np, torch
import bench
from ydorbslam_b200 import OrbExtractor
from ydorbslam_b200._lib import check, lib
B, W, H = 256, 640, 480
frames = bench.make_frames(B, 0)
ex = OrbExtractor(*bench.EXTRACTOR_ARGS, max_width=W, max_height=H, max_batch=B, device=0)
cap = ex.capacity
h_img = torch.from_numpy(frames).pin_memory()
h_kps = torch.zeros(B, cap, 28, dtype=torch.uint8).pin_memory()
h_desc = torch.zeros(B, cap, 32, dtype=torch.uint8).pin_memory()
h_n = torch.zeros(B, dtype=torch.int32).pin_memory()
def ext():
    check(lib().ydorb_extract_batch(ex._h, C.c_void_p(h_img.data_ptr()), W, H, W, C.c_size_t(W * H), B, C.c_void_p(h_kps.data_ptr()), C.c_void_p(h_desc.data_ptr()), cap, C.c_void_p(h_n.data_ptr())))
def timeit(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
for c in (4, 6, 8):
    os.environ["YDORB_HOST_CHUNKS"] = str(c)
    print("chunks", c, "extract_batch ms", round(timeit(ext), 3))
del os.environ["YDORB_HOST_CHUNKS"]
print("default plan", "extract_batch ms", round(timeit(ext), 3))
for plan in ("16,48,64", "8,24,32,64", "16,32,48,64", "24,40,64", "16,48,96", "32,64,80", "16,48,64,64,32,32", "8,16,40,64", "16,48,64,48,40,24,16"):
    os.environ["YDORB_HOST_PLAN"] = plan
    print("plan", plan, "extract_batch ms", round(timeit(ext), 3))
del os.environ["YDORB_HOST_PLAN"]
# device-resident, same handle, by batch size
d_img = torch.from_numpy(frames).cuda()
d_kps = torch.zeros(B, cap, 28, dtype=torch.uint8, device="cuda"); d_desc = torch.zeros(B, cap, 32, dtype=torch.uint8, device="cuda"); d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
for nb in (256, 128, 64, 32):
    def dev():
        for f0 in range(0, B, nb):
            ex.extract_batch_device(d_img[f0:].data_ptr(), W, H, nb, d_kps[f0:].data_ptr(), d_desc[f0:].data_ptr(), cap, d_n[f0:].data_ptr())
        ex.sync()
    print("device, sub-batch", nb, "ms per 256", round(timeit(dev), 3))
