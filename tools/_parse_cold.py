import os, sys, time, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch, gen_inputs
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
data = open(gen_inputs.ensure_synth(4032, 3024, 6510, 90), "rb").read()
sess = DecodeSession("cuda:0"); host = pinned_bytes(data); sess.parse(host); planes = sess.alloc_planes()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def med(f, n=40):
    v = []
    for i in range(n + 5):
        x = f()
        if i >= 5: v.append(x)
    return statistics.median(v)
def t_parse():
    t0 = time.perf_counter(); sess.parse(host); return (time.perf_counter() - t0) * 1e6
def A():
    flush.fill_(1); torch.cuda.synchronize()
    t = t_parse(); sess.transfer(); sess.decode(planes); torch.cuda.synchronize(); return t
def B():
    flush.fill_(1); torch.cuda.synchronize(); return t_parse()
def Cc():
    flush.fill_(1); torch.cuda.synchronize()
    t_parse(); t = t_parse(); sess.transfer(); sess.decode(planes); torch.cuda.synchronize(); return t
def D():  # no flush
    t = t_parse(); sess.transfer(); sess.decode(planes); torch.cuda.synchronize(); return t
def E():  # transfer only
    torch.cuda.synchronize(); t = t_parse(); sess.transfer(); torch.cuda.synchronize(); return t
def F():  # sleep instead of gpu work
    time.sleep(0.001); return t_parse()
for name, f in (("A full flow", A), ("B flush+parse only", B), ("C second parse in flow", Cc), ("D no flush", D), ("E parse+transfer", E), ("F sleep 1ms + parse", F), ("A again", A)):
    print(name, round(med(f), 1), "us")
