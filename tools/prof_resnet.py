"""One SE-ResNet34 backend pass on resident features for ncu (development tool): warm-up passes, then ONE pass bracketed by
cudaProfilerStart/Stop.   python tools/prof_resnet.py [--batch 64] [--frames 201]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--frames", type=int, default=201)
    a = ap.parse_args()
    import torch
    from occm_b200.models.senet import ssl_resnet34
    dev = torch.device("cuda", 0)
    m = ssl_resnet34(dev, dtype="bf16")
    feat = torch.randn(a.batch, a.frames, 1024, device=dev)
    with torch.no_grad():
        for _ in range(2):
            com, des = m.resnet34_forward(feat)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            m.resnet34_forward(feat)
        e1.record()
        torch.cuda.synchronize()
        print("SE-ResNet34 backend:", round(e0.elapsed_time(e1) / 5, 3), "ms per pass")
        torch.cuda.profiler.start()
        com, des = m.resnet34_forward(feat)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    print("finite:", bool(torch.isfinite(com).all()), float(com.abs().max()))


if __name__ == "__main__":
    main()
