"""Timeline of SequenceFrontEnd.run_host: when each pipeline stage's H2D, kernels and D2H finish (ms from start)."""
import sys, os, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
synth = bench.load_synth()
import slam_rs_b200 as S
from slam_rs_b200.frontend import SequenceFrontEnd

T, H, W = 1000, 376, 1241
hc = int(sys.argv[1]) if len(sys.argv) > 1 else 250
do_h2d = "noh2d" not in sys.argv
do_d2h = "nod2h" not in sys.argv
frames = synth.sequence(H, W, T)
fe = SequenceFrontEnd(W, H, nfeatures=2000, max_frames=T, chunk=1000, host_chunk=hc)
pin = torch.from_numpy(frames).pin_memory()
out = fe.alloc_host_outputs()
for _ in range(3):
    fe.run_host(pin, out); torch.cuda.synchronize()

# instrumented copy of run_host
cur = torch.cuda.current_stream()
def ev(stream):
    e = torch.cuda.Event(enable_timing=True); e.record(stream); return e
c = fe.host_chunk
bounds, a = [], 0
for size in (max(c // 4, 1), max(c // 2, 1)):
    if T - a > c: bounds.append((a, a + size)); a += size
while T - a > c + c // 2: bounds.append((a, a + c)); a += c
if T - a > c: bounds.append((a, T - c // 2)); a = T - c // 2
bounds.append((a, T))
torch.cuda.synchronize()
t0 = ev(cur)
for s in (fe.stream, fe.copy_stream, fe.d2h_stream): s.wait_stream(cur)
h2d, comp, d2h = [], [], []
with torch.cuda.stream(fe.copy_stream):
    for a, b in bounds:
        if do_h2d:
            fe.frames_dev[a:b].copy_(pin[a:b], non_blocking=True)
        h2d.append(ev(fe.copy_stream))
for i, (a, b) in enumerate(bounds):
    fe.stream.wait_event(h2d[i]); fe.extract(fe.frames_dev, a, b); fe.match(a, b); comp.append(ev(fe.stream))
    fe.d2h_stream.wait_stream(fe.stream)
    with torch.cuda.stream(fe.d2h_stream):
        for k in ("n", "kps", "desc", "good", "n_good"):
            if do_d2h:
                out[k][a:b].copy_(getattr(fe, k)[a:b], non_blocking=True)
        d2h.append(ev(fe.d2h_stream))
torch.cuda.synchronize()
for i, (a, b) in enumerate(bounds):
    print("stage %d frames %4d-%4d  h2d done %6.2f  kernels done %6.2f  d2h done %6.2f" %
          (i, a, b, t0.elapsed_time(h2d[i]), t0.elapsed_time(comp[i]), t0.elapsed_time(d2h[i])))
