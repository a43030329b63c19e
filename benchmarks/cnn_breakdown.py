#!/usr/bin/env python
"""Per-op device time of the CNN step on the native kernels (CUDA events, mean of N reps)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def timeit(fn, reps=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--width', type=int, default=1)
    ap.add_argument('--batch', type=int, default=512)
    a = ap.parse_args()
    import pytorch_distributed_mnist_b200 as pdm
    from pytorch_distributed_mnist_b200.ops import _lib, gemm as G, cnn_ops as C
    lib = _lib.lib()
    dev = torch.device('cuda', 0)
    B, w = a.batch, a.width
    c1, c2, H = 32 * w, 64 * w, 128 * w
    m = pdm.CNN(width=w).to(dev)
    x = torch.randn(B, 784, device=dev)
    sp = _lib.stream_ptr
    a1 = torch.empty(B * 676, c1, device=dev)
    wk = torch.empty(c2, 9 * c1, device=dev)
    wd = torch.empty(c1, 9 * c2, device=dev)
    z2 = torch.empty(B * 676, c2, device=dev)
    pooled = torch.empty(B, c2 * 144, device=dev)
    code = torch.empty(B * c2 * 144, dtype=torch.uint8, device=dev)
    h1 = torch.zeros(B, H, device=dev)
    logits = torch.empty(B, 10, device=dev)
    dl = torch.randn(B, 10, device=dev)
    dz1 = torch.randn(B, H, device=dev)
    dw2 = torch.zeros(10, H, device=dev)
    db2 = torch.zeros(10, device=dev)
    db1 = torch.zeros(H, device=dev)
    dpool = torch.randn(B, c2 * 144, device=dev)
    dy2 = torch.empty(B * 676, c2, device=dev)
    dbc2 = torch.zeros(c2, device=dev)
    dw1 = torch.zeros(c1, 9, device=dev)
    dbc1 = torch.zeros(c1, device=dev)
    gw = torch.zeros(c2, 9 * c1, device=dev)
    rows = []

    def add(name, fn, flop=0.0, bytes_=0.0):
        us = timeit(fn)
        rows.append((name, us, flop / us / 1e6 if flop else 0.0, bytes_ / us / 1e3 if bytes_ else 0.0))

    add('conv1 fwd (FMA)', lambda: lib.pdm_conv1_fwd(x.data_ptr(), 1, m.conv1.weight.data_ptr(), m.conv1.bias.data_ptr(), a1.data_ptr(), B, c1, sp()),
        2.0 * B * 676 * c1 * 9, 4.0 * B * 676 * c1)
    add('conv2 weight relayout', lambda: lib.pdm_conv_relayout(m.conv2.weight.data_ptr(), wk.data_ptr(), wd.data_ptr(), c2, c1, 0, sp()))
    add('conv2 fwd (tcgen05 implicit GEMM)', lambda: G.gemm(a1, wk, out=z2, mode=G.MODE_CONV, img_w=26, cblocks=c1 // 32, M=B * 676, N=c2, K=9 * c1),
        2.0 * B * 676 * c2 * 9 * c1, 4.0 * B * 676 * (c1 + c2))
    add('pool+bias+relu+dropout fwd', lambda: lib.pdm_pool_fwd(z2.data_ptr(), m.conv2.bias.data_ptr(), pooled.data_ptr(), code.data_ptr(), B, c2, 0.25, 7, sp()),
        0, 4.0 * B * 676 * c2)
    tiles = ((B + 127) // 128) * ((H + 127) // 128)
    splits = C._splits_for(tiles, (c2 * 144 + 31) // 32)
    add('fc1 fwd (tcgen05 split-K {})'.format(splits), lambda: (h1.zero_(), G.gemm(pooled, m.fc1.weight, out=h1, block_n=128, splits=splits)),
        2.0 * B * H * c2 * 144, 4.0 * (B + H) * c2 * 144)
    add('fc1 bias+relu', lambda: lib.pdm_bias_relu(h1.data_ptr(), m.fc1.bias.data_ptr(), B, H, sp()))
    add('fc2 fwd (dropout+linear)', lambda: lib.pdm_fcout_fwd(h1.data_ptr(), m.fc2.weight.data_ptr(), m.fc2.bias.data_ptr(), logits.data_ptr(), B, H, 0.5, 9, sp()))
    add('fc2 bwd', lambda: lib.pdm_fcout_bwd(dl.data_ptr(), h1.data_ptr(), m.fc2.weight.data_ptr(), dz1.data_ptr(), dw2.data_ptr(), db2.data_ptr(), db1.data_ptr(), B, H, 0.5, 9, sp()))
    add('fc1 wgrad (tcgen05 MN/MN)', lambda: G.gemm(dz1, pooled, a_mn_major=True, b_mn_major=True, block_n=64),
        2.0 * B * H * c2 * 144, 4.0 * H * c2 * 144)
    add('fc1 dgrad (tcgen05 K/MN)', lambda: G.gemm(dz1, m.fc1.weight, b_mn_major=True, block_n=256),
        2.0 * B * H * c2 * 144, 4.0 * (B + H) * c2 * 144)
    add('pool bwd', lambda: lib.pdm_pool_bwd(dpool.data_ptr(), code.data_ptr(), dy2.data_ptr(), dbc2.data_ptr(), B, c2, 0.25, 7, sp()),
        0, 4.0 * B * 676 * c2)
    tiles = ((c2 + 127) // 128) * ((9 * c1 + 255) // 256)
    splits = C._splits_for(tiles, (B * 676 + 31) // 32)
    add('conv2 wgrad (tcgen05 shifted, split-K {})'.format(splits), lambda: (gw.zero_(), G.gemm(dy2, a1, out=gw, a_mn_major=True, b_mn_major=True, mode=G.MODE_WGRAD, img_w=26, cblocks=c1 // 32, M=c2, N=9 * c1, K=B * 676, splits=splits)),
        2.0 * B * 676 * c2 * 9 * c1, 4.0 * B * 676 * (c1 + c2))
    add('conv2 dgrad (tcgen05 shifted)', lambda: G.gemm(dy2, wd, out=a1.clone(), mode=G.MODE_CONV, img_w=26, cblocks=c2 // 32, shift_sign=-1, M=B * 676, N=c1, K=9 * c2, mask=a1),
        2.0 * B * 676 * c2 * 9 * c1, 4.0 * B * 676 * (c1 + c2))
    add('conv1 bwd (FMA reduce)', lambda: lib.pdm_conv1_bwd(x.data_ptr(), 1, a1.data_ptr(), dw1.data_ptr(), dbc1.data_ptr(), B, c1, sp()), 2.0 * B * 676 * c1 * 10, 4.0 * B * 676 * c1)
    tot = sum(r[1] for r in rows)
    print('width {} batch {}: sum of ops {:.1f} us'.format(w, B, tot))
    for name, us, tf, gbs in rows:
        print('  {:44s} {:9.1f} us  {:7.1f} TFLOP/s  {:8.1f} GB/s'.format(name, us, tf, gbs))


if __name__ == '__main__':
    main()
