#!/usr/bin/env python
"""Cross-checks the oracle's Inception-v3 definition against torchvision on CPU and writes
tests/golden/inception_v3.npz (logits of both for two seeded images).

torchvision.models.inception_v3(weights=None, aux_logits=False, transform_input=False, init_weights=False) is an
independent statement of the architecture; the oracle's seeded weights are loaded into it (conv weight = folded
weight, BN gamma=1, beta=bias, mean=0, var=1-eps so that the BN scale is exactly 1).
"""
import os
import sys

import numpy as np
import torch
import torchvision

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle.inception as oi  # noqa: E402

SEED, CLASSES = 2024, 1000


def load_weights(model):
    for name, mod in model.named_modules():
        if isinstance(mod, torchvision.models.inception.BasicConv2d):
            cout, cin, kh, kw = mod.conv.weight.shape
            w, b = oi.conv_params(SEED, name, cout, kh, kw, cin)
            with torch.no_grad():
                mod.conv.weight.copy_(torch.from_numpy(w).permute(0, 3, 1, 2))
                mod.bn.weight.fill_(1.0)
                mod.bn.bias.copy_(torch.from_numpy(b))
                mod.bn.running_mean.zero_()
                mod.bn.running_var.fill_(1.0 - mod.bn.eps)
    w, b = oi.fc_params(SEED, CLASSES)
    with torch.no_grad():
        model.fc.weight.copy_(torch.from_numpy(w))
        model.fc.bias.copy_(torch.from_numpy(b))


def main():
    torch.manual_seed(0)
    model = torchvision.models.inception_v3(weights=None, aux_logits=False, transform_input=False, init_weights=False)
    model.eval()
    load_weights(model)
    x = np.random.default_rng(7).standard_normal((2, 299, 299, 3)).astype(np.float32)
    with torch.no_grad():
        tv = model(torch.from_numpy(x).permute(0, 3, 1, 2).contiguous()).numpy()
    fp32 = oi.Net(CLASSES, SEED, 0).forward(x)
    bf16 = oi.Net(CLASSES, SEED, 1).forward(x)
    rel = np.abs(tv - fp32).max() / np.abs(tv).max()
    print("oracle fp32 vs torchvision: max abs diff / max |logit| =", rel)
    print("oracle bf16 vs fp32       :", np.abs(bf16 - fp32).max() / np.abs(fp32).max())
    top = lambda a: np.argsort(-a, axis=1, kind="stable")[:, :5]
    print("top5 tv  ", top(tv).tolist(), "\ntop5 fp32", top(fp32).tolist(), "\ntop5 bf16", top(bf16).tolist())
    assert rel < 1e-4, rel
    assert np.array_equal(top(tv), top(fp32))
    np.savez_compressed(os.path.join(HERE, "inception_v3.npz"), image_seed=7, seed=SEED, logits_torchvision=tv,
                        logits_oracle_fp32=fp32, logits_oracle_bf16=bf16)
    print("wrote inception_v3.npz")


if __name__ == "__main__":
    main()
