"""e2e (host images in, host predictions out) for the streaming variants of MapAnything.infer."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mapanything_b200 import MapAnything, synthetic
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
model = MapAnything(device=dev, head_chunk=8); model.load_state_dict(synthetic.synthetic_state_dict())
views = [{"img": v["img"].pin_memory(), "data_norm_type": ["dinov2"]} for v in synthetic.synthetic_views(8, 392, 518, seed=1)]
host = {}
def old_way():
    preds = model.infer([dict(v) for v in views])
    for i, p in enumerate(preds):
        for k, t in p.items():
            if (i, k) not in host: host[(i, k)] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            host[(i, k)].copy_(t, non_blocking=True)
    torch.cuda.synchronize()
def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print(f"device outputs + copies afterwards: {timed(old_way):.2f} ms")
for c in (1, 2, 4, 8, (4, 2, 2), (4, 2, 1), (3, 3, 2), (5, 2, 1), (4, 3, 1), (6, 1), (2, 2, 2, 1)):
    print(f"host_outputs stream_chunk={c}: {timed(lambda: model.infer([dict(v) for v in views], host_outputs=True, stream_chunk=c)):.2f} ms")
imgs = torch.cat([v["img"] for v in views], 0).to(dev)
print(f"resident forward+post: {timed(lambda: model.run_resident(imgs, 392, 518)):.2f} ms")
t0 = time.perf_counter()
for _ in range(10): model.run_resident(imgs, 392, 518)
cpu_ms = (time.perf_counter() - t0) / 10 * 1e3; torch.cuda.synchronize()
print(f"CPU enqueue time of run_resident: {cpu_ms:.2f} ms")
ts = []
for _ in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter(); model.run_resident(imgs, 392, 518); ts.append((time.perf_counter() - t0) * 1e3)
torch.cuda.synchronize()
print("CPU enqueue time of one run_resident on an empty queue (ms):", [round(t, 2) for t in ts])
