import sys, torch
sys.path.insert(0, ".")
from camouflage_multimodal_b200 import synth, _lib
from camouflage_multimodal_b200.region_graph import canny_edges
b = synth.make_batch(0, 32, seed=1234, device=torch.device("cuda")).as_decoded()
for _ in range(3): e = canny_edges(b.images)
torch.cuda.synchronize()
_lib.timing_enable(True)
for _ in range(10): e = canny_edges(b.images)
torch.cuda.synchronize()
rep = _lib.timing_report(); _lib.timing_enable(False)
tot = 0
for k, (c, ms) in sorted(rep.items(), key=lambda kv: -kv[1][1]):
    print(k, c, round(ms / 10 * 1000, 1), "us/batch"); tot += ms / 10
print("total us", tot * 1000, "edges frac", float(e.float().mean()), "match stored canny", float((e.bool() == b.canny.bool()).float().mean()))
