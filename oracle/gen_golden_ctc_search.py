"""Golden vectors for the CTC prefix beam search (test infrastructure, not the product).

Runs the REAL reference implementation, ``ASRModel._ctc_prefix_beam_search``
(/root/reference/wenet/transformer/asr_model.py:326-409), on seeded log-probs, checks that
``oracle.reference.ctc_prefix_beam_search`` reproduces it (token lists exactly, scores to 1e-12)
and writes ``tests/golden/ctc_search.npz``.  Run in the build container only
(``python oracle/gen_golden_ctc_search.py``); the fixture travels, /root/reference does not.
"""
from __future__ import annotations

import os
import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")
GOLD = ROOT / "tests" / "golden"
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True


class _FakeCtc:
    @staticmethod
    def log_softmax(x):
        return torch.log_softmax(x, dim=2)


class _FakeSelf:
    """what _ctc_prefix_beam_search reads from the model: _forward_encoder and ctc.log_softmax;
    the 'speech' fed in is already the CTC head's logits."""
    ctc = _FakeCtc()

    @staticmethod
    def _forward_encoder(speech, speech_lengths, *a, **k):
        return speech, None


def cases():
    g = torch.Generator().manual_seed(99)
    out = []
    # (name, logits [T,V], beam)
    out.append(("flat", torch.randn(40, 50, generator=g), 5))
    x = torch.randn(60, 30, generator=g) * 3.0
    x[:, 0] += 4.0  # blank-dominated, peaky: long runs of blank / repeats
    out.append(("peaky", x, 4))
    out.append(("tiny_vocab", torch.randn(25, 6, generator=g) * 1.5, 4))  # heavy prefix merging
    y = torch.randn(30, 12, generator=g)
    y[:, 3] += 2.5  # one dominant symbol: repeat handling (*ss -> *s, *s-s -> *ss)
    out.append(("repeats", y, 6))
    out.append(("beam1", torch.randn(20, 9, generator=g), 1))
    out.append(("beam16", torch.randn(35, 40, generator=g) * 2.0, 16))
    return out


def main():
    stub = types.ModuleType("turtle")
    stub.forward = None
    sys.modules.setdefault("turtle", stub)
    sys.path.insert(0, str(REF))
    from wenet.transformer.asr_model import ASRModel
    from oracle import reference as orc

    blob = {}
    names = []
    for name, logits, beam in cases():
        speech = logits.unsqueeze(0)
        hyps, _ = ASRModel._ctc_prefix_beam_search(_FakeSelf(), speech, torch.tensor([logits.size(0)]), beam)
        mine = orc.ctc_prefix_beam_search(torch.log_softmax(speech, dim=2)[0], beam)
        assert [h[0] for h in hyps] == [h[0] for h in mine], name
        assert np.allclose([h[1] for h in hyps], [h[1] for h in mine], rtol=1e-12, atol=0), name
        T = logits.size(0)
        toks = np.full((beam, T), -1, dtype=np.int64)
        for k, (prefix, _) in enumerate(hyps):
            toks[k, :len(prefix)] = prefix
        blob[f"{name}.logits"] = logits.numpy()
        blob[f"{name}.beam"] = np.int64(beam)
        blob[f"{name}.n"] = np.int64(len(hyps))
        blob[f"{name}.tokens"] = toks
        blob[f"{name}.lens"] = np.array([len(h[0]) for h in hyps], dtype=np.int64)
        blob[f"{name}.scores"] = np.array([h[1] for h in hyps], dtype=np.float64)
        names.append(name)
        print(f"{name}: T={T} V={logits.size(1)} beam={beam} best={hyps[0][0][:8]}... score={hyps[0][1]:.6f}")
    blob["names"] = np.array(names)
    # ---- forced alignment: the real forced_align (wenet/utils/ctc_util.py:27-83)
    from wenet.utils.ctc_util import forced_align as ref_align
    g = torch.Generator().manual_seed(123)
    anames = []
    for name, T, V, y in [("plain", 30, 20, [3, 7, 7, 2, 9]), ("tight", 12, 15, [4, 4, 5, 6, 6, 1]),
                          ("single", 9, 8, [5]), ("infeasible", 5, 10, [1, 2, 3, 4]),
                          ("long", 120, 40, [int(v) for v in torch.randint(1, 40, (30,), generator=g)]),
                          ("one_frame", 1, 6, [2, 3])]:
        lp = torch.log_softmax(torch.randn(T, V, generator=g) * 2.0, dim=1)
        yt = torch.tensor(y, dtype=torch.int64)
        want = [int(v) for v in ref_align(lp, yt, 0)]
        mine = orc.forced_align(lp, yt, 0)
        assert want == mine, (name, want, mine)
        blob[f"align.{name}.logp"] = lp.numpy()
        blob[f"align.{name}.y"] = yt.numpy()
        blob[f"align.{name}.alignment"] = np.array(want, dtype=np.int64)
        anames.append(name)
        print(f"align {name}: T={T} L={len(y)} -> {want[:16]}")
    blob["align_names"] = np.array(anames)
    GOLD.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(GOLD / "ctc_search.npz", **blob)
    print("written", GOLD / "ctc_search.npz", (GOLD / "ctc_search.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    main()
