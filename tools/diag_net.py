"""GPU diagnostic: per-node error of the B200 ResNet34 vs (a) bf16 emulation and (b) fp32 torch."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import torch
from gpu_net_helpers import bf16_round, emulate_ir
from oracle.forward import forward_probs, make_resnet34
from wsinsight_b200.modellib.engine import B200Network

def padded(x):
    B, _, S, _ = x.shape
    buf = torch.zeros((B, S + 6, S + 6, 4), dtype=torch.bfloat16, device="cuda")
    buf[:, 3:3 + S, 3:3 + S, :3] = x.permute(0, 2, 3, 1).to(torch.bfloat16).cuda()
    return buf

for (K, seed, B, maxb) in [(1, 3, 5, 8), (2, 0, 16, 16), (1, 3, 8, 8)]:
    m = make_resnet34(K, seed)
    g = torch.Generator().manual_seed(2)
    x = bf16_round(torch.randn((B, 3, 224, 224), generator=g) * 1.5)
    net = B200Network.from_torchscript(torch.jit.script(m), K, input_size=224, max_batch=maxb)
    p = net(net.pack_input(x), B).cpu()
    acts, pe = emulate_ir(net.nodes, x)
    pf = forward_probs(m, x)
    print(f"K={K} seed={seed} B={B} max_batch={maxb}")
    print("  gpu   ", p.flatten()[:8].tolist())
    print("  emul  ", pe.flatten()[:8].tolist())
    print("  fp32  ", pf.flatten()[:8].tolist())
    print(f"  max|gpu-emul|={(p-pe).abs().max():.3e}  max|gpu-fp32|={(p-pf).abs().max():.3e}  max|emul-fp32|={(pe-pf).abs().max():.3e}")
    for i, n in enumerate(net.nodes):
        if n.kind in ("conv", "pool"):
            a = net.activation(i, B).float().permute(0, 3, 1, 2).cpu()
            a = a[:, net.channel_offset(i):net.channel_offset(i) + n.shape[0]]
            e = acts[i]
            d = (a - e).abs()
            rel = (d.mean() / (e.abs().mean() + 1e-9)).item()
            print(f"  node {i:2d} {n.kind:4s} {tuple(n.shape)} rel_mean_err={rel:.2e} max_err={d.max().item():.3e} per-image max {[round(v,3) for v in d.flatten(1).max(1).values.tolist()]}")
    net.close()
