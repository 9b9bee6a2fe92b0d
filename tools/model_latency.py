"""Developer tool: latency / throughput of the full ITAViTLSTM forward (uint8 frames resident on the GPU) vs batch, including the
batch-1 case the deployed loop (samples/inference_udp_FPGA_custom_dispatch/main.cpp:150-231) runs; and the same end to end from
pinned host packets.   python tools/model_latency.py [--out profiles/r2_model_latency.txt]"""
import argparse, os, statistics, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
ap = argparse.ArgumentParser(); ap.add_argument("--out", default=""); args = ap.parse_args()
dev = torch.device("cuda", 0)
m, _ = bench._load_model(dev)
g = np.random.default_rng(1)
import importlib
M = importlib.import_module("drone-oa-iree-vit-accelerator_b200.model")
lines = ["batch | device latency us (median of 30) | device k frames/s | host-to-host latency us | host-to-host k frames/s | CUDA-graph step: device latency us | host-to-host latency us"]
for B in (1, 2, 8, 32, 128, 512, 2048, 4096, 16384):
    img_h = torch.from_numpy(g.integers(0, 256, (B, 60, 90), dtype=np.uint8)).pin_memory()
    vel_h = torch.from_numpy(g.random((B, 1), dtype=np.float32)).pin_memory()
    quat_h = torch.from_numpy(g.random((B, 4), dtype=np.float32)).pin_memory()
    out_h = torch.empty((B, 3), dtype=torch.float32).pin_memory()
    img, vel, quat = img_h.to(dev), vel_h.to(dev), quat_h.to(dev)
    hid = None
    for _ in range(5):
        out, hid = m([img, vel, quat] + ([hid] if hid is not None else []))
    torch.cuda.synchronize()
    lat = []
    for _ in range(30):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out, hid = m([img, vel, quat, hid])
        e1.record(); e1.synchronize()
        lat.append(e0.elapsed_time(e1) * 1e3)
    h2h = []
    for _ in range(30):
        t0 = time.perf_counter()
        out, hid = m([img_h.to(dev, non_blocking=True), vel_h.to(dev, non_blocking=True), quat_h.to(dev, non_blocking=True), hid])
        out_h.copy_(out, non_blocking=True)
        torch.cuda.synchronize()
        h2h.append((time.perf_counter() - t0) * 1e6)
    l, h = statistics.median(lat), statistics.median(h2h)
    step = M.GraphedStep(m, B, dev)
    for _ in range(3):
        step(img, vel, quat)
    torch.cuda.synchronize()
    glat, gh2h = [], []
    for _ in range(30):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step(img, vel, quat)
        e1.record(); e1.synchronize()
        glat.append(e0.elapsed_time(e1) * 1e3)
    for _ in range(30):
        t0 = time.perf_counter()
        o = step(img_h, vel_h, quat_h)
        out_h.copy_(o, non_blocking=True)
        torch.cuda.synchronize()
        gh2h.append((time.perf_counter() - t0) * 1e6)
    # the graphed step computes what the eager step computes (same state trajectory from zero)
    step.reset_state()
    o1 = step(img, vel, quat).clone()
    o_ref, _ = m([img, vel, quat])
    assert torch.allclose(o1, o_ref, atol=1e-6), (o1 - o_ref).abs().max().item()
    lines.append(f"{B:5d} | {l:10.1f} | {B / l * 1e3:10.1f} | {h:10.1f} | {B / h * 1e3:10.1f} | {statistics.median(glat):10.1f} | {statistics.median(gh2h):10.1f}")
    del step
txt = "\n".join(lines)
print(txt)
if args.out:
    open(args.out, "w").write("# full ITAViTLSTM forward, one B200, LSTM state carried between calls (tools/model_latency.py)\n" + txt + "\n")
