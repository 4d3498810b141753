#!/usr/bin/env python
"""Diagnostic: accuracy of kernel (c) (Cholesky log-det forward + adjoint) and the Gram adjoint on an
ill-conditioned metric (cond ~ 1e4, like MNIST-shape d = 64) against fp64, next to torch's own fp32 path."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cmf_b200 import kernels as K
torch.manual_seed(0)
B, d, n = 16, 64, 784
for cond in (1e2, 5e3, 3e4):
    V = torch.randn(B, d, n, device="cuda")
    scale = torch.logspace(0, -0.5 * torch.log10(torch.tensor(cond)).item(), d, device="cuda")
    mix = torch.linalg.qr(torch.randn(d, d, device="cuda"))[0]
    V = mix @ (scale.view(1, d, 1) * V)
    G = K.gram_fwd(V)
    G64 = G.double()
    c = float(torch.linalg.cond(G64).max())
    Ginv64 = torch.linalg.inv(G64)
    ld64 = torch.linalg.slogdet(G64)[1]
    L, logdet, l1o, l1d, info = K.chol_logdet_fwd(G)
    gG = K.chol_logdet_bwd(L, G, torch.ones(B, device="cuda"), None, "none")
    rel = lambda a, b: float(((a.double() - b).flatten(1).norm(dim=1) / b.flatten(1).norm(dim=1)).max())
    # torch fp32
    Gt = G.clone().requires_grad_(True)
    Lt = torch.linalg.cholesky(Gt)
    ldt = 2 * torch.log(torch.diagonal(Lt, dim1=-2, dim2=-1)).sum(1)
    ldt.sum().backward()
    gt = 0.5 * (Gt.grad + Gt.grad.transpose(1, 2))
    # fp32 via explicit inverse
    gi = torch.linalg.inv(G)
    print(f"cond {c:.2e}: logdet err kernel {float((logdet.double()-ld64).abs().max()):.2e} torch {float((ldt.double()-ld64).abs().max()):.2e} | "
          f"G^-1 rel-F err: kernel {rel(gG, Ginv64):.2e}  torch chol-backward {rel(gt, Ginv64):.2e}  torch.inv fp32 {rel(gi, Ginv64):.2e} | "
          f"L err kernel {rel(L, torch.linalg.cholesky(G64)):.2e} torch {rel(Lt.detach(), torch.linalg.cholesky(G64)):.2e}")
    S = Ginv64
    gV64 = (S + S.transpose(1, 2)) @ V.double()
    print(f"    gram adjoint (given exact S in fp32): kernel {rel(K.gram_bwd(V, S.float().contiguous()), gV64):.2e}  torch bmm {rel((S.float()+S.float().transpose(1,2)) @ V, gV64):.2e}")
    print(f"    gram forward vs fp64: kernel tc {rel(G, V.double() @ V.double().transpose(1,2)):.2e} fma {rel(K.gram_fwd(V, path='fma'), V.double() @ V.double().transpose(1,2)):.2e} torch bmm {rel(V @ V.transpose(1,2), V.double() @ V.double().transpose(1,2)):.2e}")
