import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")]
import numpy as np, torch
import inputs
import jpeggpu_b200 as jb
from jpeggpu_b200.batch import FusedBatch, make_items
lib = jb.load_library()
good = inputs.pillow_jpeg(w=640, h=480, seed=71, mode="YCbCr", subsampling=2)
for files in ([good, good], [good, good[:200], good]):
    hosts = []
    for f in files:
        t = torch.empty(len(f), dtype=torch.uint8, pin_memory=True)
        t.numpy()[:] = np.frombuffer(f, dtype=np.uint8)
        hosts.append(t)
    items = make_items([(t.data_ptr(), t.numel()) for t in hosts])
    print("ptrs", [hex(t.data_ptr()) for t in hosts])
    fb = FusedBatch(lib)
    need, st = fb.plan(items, strict=False)
    print("plan", need, st, [it.status for it in items])
    arena = torch.zeros(need, dtype=torch.uint8, device="cuda:0")
    planes = {}
    for i, it in enumerate(items):
        if it.status: continue
        ps = [torch.zeros((it.info.sizes_y[c], it.info.sizes_x[c]), dtype=torch.uint8, device="cuda:0") for c in range(it.info.num_components)]
        for c, p in enumerate(ps):
            it.img.image[c] = p.data_ptr(); it.img.pitch[c] = p.stride(0)
        planes[i] = ps
    st = fb.decode(items, arena.data_ptr(), arena.numel(), torch.cuda.current_stream().cuda_stream, strict=False)
    torch.cuda.synchronize()
    print("decode status", st)
    for i in planes:
        si = fb.read_scan_info(i, arena.data_ptr(), torch.cuda.current_stream().cuda_stream)
        print(i, si.num_subsequences, si.num_segments, si.scan_end, si.status, int(planes[i][0].sum()))
    fb.close()
