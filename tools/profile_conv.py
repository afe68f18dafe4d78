#!/usr/bin/env python
"""Per-layer timing of the tensor-core convolution (K3) on the Inception-v3 shapes at batch 64.

  python tools/profile_conv.py [--batch 64] [--iters 20] [--only NAME[,NAME]] [--json out.json]

Each layer runs through ftf_conv2d_bf16_dev on torch-owned buffers; time = CUDA events over `iters` back-to-back
launches after 3 warm-ups (an L2 flush between layers, not between iterations: weights stay L2-resident as they do
inside the network, activations of the big layers exceed L2 anyway).  Prints algorithmic TFLOP/s and the fraction of the
measured bf16 peak.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# name, H, W, Cin, Cout, KH, KW, stride, ph, pw   (Cout of fused 1x1 groups = sum of the siblings)
LAYERS = [
    ("Conv2d_2a_3x3", 149, 149, 32, 32, 3, 3, 1, 0, 0),
    ("Conv2d_2b_3x3", 147, 147, 32, 64, 3, 3, 1, 1, 1),
    ("Conv2d_3b_1x1", 73, 73, 64, 80, 1, 1, 1, 0, 0),
    ("Conv2d_4a_3x3", 73, 73, 80, 192, 3, 3, 1, 0, 0),
    ("Mixed_5b.1x1x4", 35, 35, 192, 208, 1, 1, 1, 0, 0),
    ("Mixed_5b.5x5", 35, 35, 48, 64, 5, 5, 1, 2, 2),
    ("Mixed_5b.3x3dbl_2", 35, 35, 64, 96, 3, 3, 1, 1, 1),
    ("Mixed_5b.3x3dbl_3", 35, 35, 96, 96, 3, 3, 1, 1, 1),
    ("Mixed_5c.1x1x4", 35, 35, 256, 240, 1, 1, 1, 0, 0),
    ("Mixed_6a.3x3s2", 35, 35, 288, 384, 3, 3, 2, 0, 0),
    ("Mixed_6a.3x3dbl_3s2", 35, 35, 96, 96, 3, 3, 2, 0, 0),
    ("Mixed_6b.1x1x4", 17, 17, 768, 640, 1, 1, 1, 0, 0),
    ("Mixed_6b.1x7_128", 17, 17, 128, 128, 1, 7, 1, 0, 3),
    ("Mixed_6b.7x1_128_192", 17, 17, 128, 192, 7, 1, 1, 3, 0),
    ("Mixed_6c.1x1x4", 17, 17, 768, 704, 1, 1, 1, 0, 0),
    ("Mixed_6c.1x7_160", 17, 17, 160, 160, 1, 7, 1, 0, 3),
    ("Mixed_6e.1x1x4", 17, 17, 768, 768, 1, 1, 1, 0, 0),
    ("Mixed_6e.7x1_192", 17, 17, 192, 192, 7, 1, 1, 3, 0),
    ("Mixed_7a.1x1x2", 17, 17, 768, 384, 1, 1, 1, 0, 0),
    ("Mixed_7a.3x3s2_320", 17, 17, 192, 320, 3, 3, 2, 0, 0),
    ("Mixed_7b.1x1x4", 8, 8, 1280, 1344, 1, 1, 1, 0, 0),
    ("Mixed_7b.1x3_384", 8, 8, 384, 384, 1, 3, 1, 0, 1),
    ("Mixed_7b.3x3_448_384", 8, 8, 448, 384, 3, 3, 1, 1, 1),
    ("Mixed_7c.1x1x4", 8, 8, 2048, 1344, 1, 1, 1, 0, 0),
    ("fc", 1, 1, 2048, 1008, 1, 1, 1, 0, 0),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="")
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    import torch
    import ftf_loader
    ftf_loader.load()
    from flink_tensorflow_b200 import device
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        peak = 1590.0
    only = set(filter(None, args.only.split(",")))
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    g = torch.Generator(device="cuda").manual_seed(1)
    for name, h, w, cin, cout, kh, kw, s, ph, pw in LAYERS:
        if only and name not in only:
            continue
        x = torch.randn((args.batch, h, w, cin), device="cuda", generator=g).to(torch.bfloat16)
        wt = (torch.randn((cout, kh, kw, cin), device="cuda", generator=g) * (2.0 / (kh * kw * cin)) ** 0.5).to(torch.bfloat16)
        b = torch.rand(cout, device="cuda", generator=g) * 0.2 - 0.1
        oh, ow = (h + 2 * ph - kh) // s + 1, (w + 2 * pw - kw) // s + 1
        out = torch.empty((args.batch, oh, ow, cout), dtype=torch.bfloat16, device="cuda")
        for _ in range(3):
            device.conv2d_bf16(x, wt, b, s, (ph, pw), out=out)
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            device.conv2d_bf16(x, wt, b, s, (ph, pw), out=out)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / args.iters
        flop = 2.0 * args.batch * oh * ow * cout * kh * kw * cin
        byts = x.numel() * 2 + wt.numel() * 2 + out.numel() * 2
        tf = flop / us / 1e6
        if os.environ.get("FTF_CONV_PROFILE") == "1":
            import ctypes as C
            from flink_tensorflow_b200 import _abi
            buf = (C.c_uint64 * 16)()
            _abi.load_library().ftf_debug_conv_profile(buf, 1)
            ctas, launches = max(1, buf[0]), args.iters + 3
            f = lambda i: buf[i] / ctas
            print(f"   per CTA-launch cycles: producer {f(1):9.0f} (wait free stage {f(2):9.0f}) | mma {f(3):9.0f} (wait operands {f(4):9.0f}, "
                  f"wait acc {f(5):8.0f}) | epilogue {f(6):9.0f} (wait acc {f(7):9.0f}) | tiles/CTA {buf[8] / ctas:.1f}")
        rows.append(dict(layer=name, us=round(us, 2), gflop=round(flop / 1e9, 2), tflops=round(tf, 1), frac_peak=round(tf / peak, 3),
                         hbm_gbs=round(byts / us / 1e3, 1), m_tiles=(args.batch * oh * ow + 127) // 128))
        print(f"{name:24s} {us:9.2f} us  {flop / 1e9:8.2f} GFLOP  {tf:8.1f} TFLOP/s  {tf / peak:6.3f} of peak  "
              f"{byts / us / 1e3:8.1f} GB/s min-traffic  m_tiles={rows[-1]['m_tiles']}")
        del x, wt, out
    tot_us = sum(r["us"] for r in rows)
    tot_fl = sum(r["gflop"] for r in rows)
    print(f"sum {tot_us:.1f} us, {tot_fl:.1f} GFLOP, {tot_fl / tot_us * 1e-3 * 1e3:.1f} TFLOP/s")
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
