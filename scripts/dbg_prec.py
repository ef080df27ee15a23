"""Where does the fused backward lose precision?  One utterance of the cfg3 shape on the GPU:
fp32 materialised chain vs a torch emulation of the bf16 operand roundings vs the fused kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import wenet_celoss_b200 as pkg
from wenet_celoss_b200 import joint as jm, _lib
from wenet_celoss_b200._lib import i32, ptr
from wenet_celoss_b200.rnnt import rnnt_grad_coef, rnnt_lattice

torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda:0"
B, T, U, V, E, J = 1, int(os.environ.get("T", 500)), int(os.environ.get("U", 100)), 5000, 256, 512
torch.manual_seed(13)
g = torch.Generator().manual_seed(13)
enc_ffn, pred_ffn, ffn_out = torch.nn.Linear(E, J), torch.nn.Linear(E, J), torch.nn.Linear(J, V)
with torch.no_grad():
    ffn_out.weight.mul_(2.0)
enc = torch.randn(B, T, E, generator=g); pred = torch.randn(B, U + 1, E, generator=g)
tgt = torch.randint(1, V, (B, U), generator=g, dtype=torch.int32).to(dev)
tl = torch.full((B,), T, dtype=torch.int32, device=dev); ul = torch.full((B,), U, dtype=torch.int32, device=dev)
e = enc_ffn(enc).detach().to(dev); p = pred_ffn(pred).detach().to(dev)
W = ffn_out.weight.detach().to(dev); b = ffn_out.bias.detach().to(dev)


def chain(hround, ground, wround):
    Hx = torch.tanh(e[:, :, None] + p[:, None])
    Hm = Hx.bfloat16().float() if hround else Hx
    Wm = W.bfloat16().float() if wround else W
    logits = (Hm @ Wm.t() + b).requires_grad_(True)
    loss = pkg.rnnt_loss(logits, tgt, tl, ul, blank=0, reduction="sum")
    loss.backward()
    G = logits.grad
    Gm = G.bfloat16().float() if ground else G
    dH = Gm @ Wm
    dz = dH * (1 - Hm * Hm)
    return dz.sum(2), dz.sum(1), loss.item(), G, Hm


def rel(a, b_):
    return ((a.double() - b_.double()).norm() / b_.double().norm()).item()


de_ref, dp_ref, l_ref, G_ref, _ = chain(False, False, False)
de_emu, dp_emu, l_emu, G_emu, H_emu = chain(True, True, True)
print(f"emulated bf16 vs fp32 : de {rel(de_emu, de_ref):.3e} dp {rel(dp_emu, dp_ref):.3e} loss {abs(l_emu-l_ref)/l_ref:.2e}")

for force in (False, True):
    jm._FORCE_LIBRARY_DH = force
    e1 = e.clone().requires_grad_(True); p1 = p.clone().requires_grad_(True)
    W1 = W.clone().requires_grad_(True); b1 = b.clone().requires_grad_(True)
    costs = jm._FusedJointLossFn.apply(e1, p1, W1, b1, tgt, tl, ul, 0, 0, True)
    costs.sum().backward()
    tag = "library dH+dz" if force else "joint_dh_kernel"
    print(f"fused [{tag:15s}] vs fp32 : de {rel(e1.grad, de_ref):.3e} dp {rel(p1.grad, dp_ref):.3e} "
          f"loss {abs(costs.sum().item()-l_ref)/l_ref:.2e} | vs emu: de {rel(e1.grad, de_emu):.3e} dp {rel(p1.grad, dp_emu):.3e}")
    dW_ref = (G_ref.reshape(-1, V).t() @ torch.tanh(e[:, :, None] + p[:, None]).reshape(-1, J))
    print(f"      dW vs fp32 {rel(W1.grad, dW_ref):.3e}  db vs fp32 {rel(b1.grad, G_ref.reshape(-1, V).sum(0)):.3e}")

# G of the GRAD kernel itself against the emulated (unrounded) G on the same bf16 operands
U1 = U + 1
w_bf16 = W.to(torch.bfloat16).contiguous(); bias_pad = jm._pad_bias(b, V)
n = B * T * U1
Hs = torch.empty(n, J, dtype=torch.bfloat16, device=dev)
lse, lpb, lpe = jm.joint_logprobs(e.contiguous(), p.contiguous(), w_bf16, bias_pad, tgt, tl, ul, V, 0, H=Hs)
alpha, beta, costs, doff = rnnt_lattice(lpb, lpe, tl, ul)
dy = torch.ones(B, device=dev)
coef = rnnt_grad_coef(lse, alpha, beta, doff, dy, tgt, tl, ul, 0)
ldg = (V + 15) // 16 * 16
G = torch.empty(n, ldg, dtype=torch.bfloat16, device=dev)
dbp = torch.zeros(16, (V + 127) // 128 * 128, device=dev)
_lib.call("b200_joint_grad_chunk_h", ptr(e), ptr(p), ptr(w_bf16), ptr(bias_pad), ptr(tgt), ptr(coef), i32(B), i32(T),
          i32(U1), i32(V), i32(J), i32(0), jm._i64(0), jm._i64(n), ptr(G), i32(ldg), ptr(Hs), i32(J), ptr(dbp), i32(0),
          _lib.stream_handle(torch.device(dev)))
torch.cuda.synchronize()
Gm = G[:, :V].float().view(B, T, U1, V)
print(f"G kernel vs emulated G (unrounded): {rel(Gm, G_emu):.3e}   bf16(G_emu) vs G_emu: {rel(G_emu.bfloat16().float(), G_emu):.3e}")
print(f"H kernel vs bf16(tanh): {rel(Hs.float().view(B, T, U1, J), H_emu):.3e}")
# per-row check of the worst rows
d = (Gm - G_emu).flatten(0, 2).norm(dim=1) / G_emu.flatten(0, 2).norm(dim=1).clamp_min(1e-20)
print("worst G rows (rel):", d.topk(5).values.tolist(), "median", d.median().item())
# alpha/beta sanity: sum_v G = 0 per row?
print("max |sum_v G_emu|", G_emu.sum(-1).abs().max().item(), "max |sum_v G_kernel|", Gm.sum(-1).abs().max().item())
