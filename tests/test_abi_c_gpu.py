"""A plain-C caller of the C ABI (no Python / torch in the calling process): tests/abi_c/greedy_caller.c
includes include/b200_transducer.h, links libb200_transducer.so and decodes a batch with
b200_rnnt_greedy - the call a serving runtime would make where the reference's C++ runtime invokes the
TorchScript greedy search (runtime/core/decoder/torch_asr_model.cc:300-319).  The token lists must be
identical to the fixture recorded from the reference's Python greedy search."""
import os
import shutil
import struct
import subprocess

import numpy as np
import pytest
import torch

from conftest import ROOT, load_oracle_model

SRC = ROOT / "tests" / "abi_c" / "greedy_caller.c"
LIB = ROOT / "wenet_celoss_b200" / "libb200_transducer.so"


def _cuda_home():
    nv = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return os.path.dirname(os.path.dirname(os.path.realpath(nv)))


def _build(tmp_path):
    exe = tmp_path / "greedy_caller"
    cuda = _cuda_home()
    cmd = ["gcc", "-O1", "-std=c99", "-Wall", "-I", str(ROOT / "include"), "-I", f"{cuda}/include", str(SRC),
           str(LIB), f"-L{cuda}/lib64", "-lcudart", f"-Wl,-rpath,{LIB.parent}", f"-Wl,-rpath,{cuda}/lib64",
           "-o", str(exe)]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    return exe


def test_c_caller_compiles_against_header(tmp_path):
    """no GPU needed: the header is valid C99 and every symbol the caller uses links"""
    if not LIB.exists():
        pytest.skip("library not built")
    _build(tmp_path)


@pytest.mark.gpu
def test_c_caller_decodes_fixture(tmp_path, golden):
    from wenet_celoss_b200.search import _enc_proj, _model_tensors
    d = golden("search_mixed")
    m = load_oracle_model(d).to("cuda:0")
    enc = torch.from_numpy(d["enc"]).to("cuda:0")
    lens = torch.from_numpy(d["lens"]).to(torch.int32)
    n_steps = int(d["n_steps"])
    B, T, _ = enc.shape
    mt = _model_tensors(m)
    enc_proj = _enc_proj(m, enc)
    max_tokens = T * n_steps
    fx = tmp_path / "fixture.bin"
    with open(fx, "wb") as f:
        f.write(struct.pack("<10i", B, T, mt["V"], mt["Emb"], mt["H"], mt["L"], mt["J"], 0, n_steps, max_tokens))
        f.write(enc_proj.detach().cpu().numpy().astype("<f4").tobytes())
        f.write(lens.numpy().astype("<i4").tobytes())
        for t in [mt["embed"], *mt["lstm"], mt["w_c"], mt["b_c"], mt["w_out"], mt["b_out"]]:
            f.write(t.detach().cpu().numpy().astype("<f4").tobytes())
    exe = _build(tmp_path)
    out = tmp_path / "out.bin"
    r = subprocess.run([str(exe), str(fx), str(out)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    raw = np.fromfile(out, dtype="<i4")
    rounds, n_tok, toks = int(raw[0]), raw[1:1 + B], raw[1 + B:].reshape(B, max_tokens)
    assert rounds > 0
    for b in range(B):
        want = [int(x) for x in d["greedy_hyps"][b] if x >= 0]
        assert toks[b, :n_tok[b]].tolist() == want, b
