"""Check (and time) the 2-CTA dW kernel against torch.mm on the same bf16 operands."""
import sys; sys.path.insert(0, '.')
import torch
from wenet_celoss_b200 import joint as jm
dev = torch.device('cuda:0')
torch.manual_seed(0)
ok = True
for (n, V, J, pad) in [(64, 256, 128, 0), (1000, 300, 128, 4), (777, 70, 256, 2), (5000, 1000, 384, 8),
                       (50500, 5000, 512, 8), (3 * 50500, 5000, 512, 8)]:
    ldg = V + pad
    G = (torch.randn(n, ldg, device=dev) * 0.1).to(torch.bfloat16)
    H = torch.tanh(torch.randn(n, J, device=dev)).to(torch.bfloat16)
    dW = torch.zeros(V, J, device=dev)
    jm.joint_dw(G, H, n, V, J, dW)
    torch.cuda.synchronize()
    ref = G[:, :V].float().t() @ H.float()
    err = (dW - ref).abs().max().item()
    scale = ref.abs().max().item()
    good = err <= 2e-3 * scale + 1e-4
    ok &= good
    print(f"n={n} V={V} J={J} ldg={ldg}: max err {err:.3e} (scale {scale:.3e}) {'ok' if good else 'MISMATCH'}", flush=True)
    if not good:
        bad = ((dW - ref).abs() > 2e-3 * scale + 1e-4)
        rows = bad.any(1).nonzero().flatten(); cols = bad.any(0).nonzero().flatten()
        print("  bad rows", rows[:8].tolist(), "...", rows[-4:].tolist(), "count", rows.numel())
        print("  bad cols", cols[:8].tolist(), "...", cols[-4:].tolist(), "count", cols.numel())
        print("  dW[0,:8]", dW[0, :8].tolist()); print("  ref[0,:8]", ref[0, :8].tolist())
if len(sys.argv) > 1 and ok:
    nb = int(sys.argv[1])
    n, V, J, ldg = nb * 50500, 5000, 512, 5008
    G = (torch.randn(n, ldg, device=dev) * 0.01).to(torch.bfloat16)
    H = torch.randn(n, J, device=dev).to(torch.bfloat16)
    dW = torch.zeros(V, J, device=dev)
    for name, f in (("joint_dw_kernel", lambda: jm.joint_dw(G, H, n, V, J, dW)),
                    ("cuBLAS G^T H", lambda: torch.mm(G[:, :V].t(), H, out_dtype=torch.float32)),
                    ("cuBLAS H^T G", lambda: torch.mm(H.t(), G, out_dtype=torch.float32))):
        for _ in range(2): f()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        for _ in range(5): f()
        b.record(); b.synchronize()
        ms = a.elapsed_time(b) / 5
        print(f"nb={nb} {name:18s} {ms:7.3f} ms  {2.0 * n * V * J / ms / 1e9:7.1f} TF/s", flush=True)
print("ALL OK" if ok else "FAILED")
