"""cuBLAS timing of the two backward GEMM shapes of the fused joint (per chunk of nb utterances)."""
import sys, torch
dev = 'cuda:0'
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n, V, J, ldg, ldh = nb * 50500, 5000, 512, 5008, 520
G = (torch.randn(n, ldg, device=dev) * 0.01).to(torch.bfloat16)
H = torch.randn(n, ldh, device=dev).to(torch.bfloat16)
W = torch.randn(V, J, device=dev).to(torch.bfloat16)
Gv = G[:, :V]
def tm(f, name, flop):
    for _ in range(2): f()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(5): f()
    b.record(); b.synchronize()
    ms = a.elapsed_time(b) / 5
    print(f"nb={nb} {name:44s} {ms:7.3f} ms  {flop / ms / 1e9:7.1f} TF/s")
f = 2.0 * n * V
tm(lambda: torch.mm(Gv, W, out_dtype=torch.float32), "dH = G W              [n,512] f32", f * 512)
tm(lambda: torch.mm(Gv, W), "dH = G W              [n,512] bf16", f * 512)
tm(lambda: torch.mm(Gv.t(), H, out_dtype=torch.float32), "dW = G^T H_aug        [V,520] f32", f * 520)
tm(lambda: torch.mm(Gv.t(), H[:, :J], out_dtype=torch.float32), "dW = G^T H            [V,512] f32", f * 512)
tm(lambda: torch.mm(H.t(), Gv, out_dtype=torch.float32), "dW^T = H_aug^T G      [520,V] f32", f * 520)
tm(lambda: torch.mm(H[:, :J].t(), Gv, out_dtype=torch.float32), "dW^T = H^T G          [512,V] f32", f * 512)
tm(lambda: torch.mm(H[:, :J].t(), G, out_dtype=torch.float32), "dW^T = H^T G(ldg)     [512,5008] f32", f * 512)
Hc = H[:, :J].contiguous()
tm(lambda: torch.mm(Hc.t(), G, out_dtype=torch.float32), "dW^T = Hc^T G(ldg)    [512,5008] f32", f * 512)
