"""Golden vectors for the hot-word CE loss (test infrastructure, not the product).

Runs the REAL ``Transducer.forward`` (/root/reference/wenet/transducer/transducer.py:79-270) with a
real ``ContextBias`` and ``hw_weight = 0.4`` in the three ``loss_mode`` branches, captures what
``self.hw_criterion`` receives (by wrapping it), checks ``oracle.reference.hw_ce_loss`` (numpy fp64)
against it and writes ``tests/golden/hw_loss.npz``.  Run in the build container only.
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


def main():
    stub = types.ModuleType("turtle")
    stub.forward = None
    sys.modules.setdefault("turtle", stub)
    sys.path.insert(0, str(REF))
    import wenet.transducer.transducer as tr
    from wenet.transducer.joint import TransducerJoint
    from wenet.transducer.predictor import RNNPredictor
    from wenet.transformer.context_bias import ContextBias
    from wenet.transformer.ctc import CTC
    from oracle import reference as orc
    from oracle.gen_golden import PassThroughEncoder

    V, E, J, H, L = 41, 32, 48, 32, 2
    B, T, U = 4, 20, 7
    blob = {}
    for mode in ("pred", "both", "sep"):
        torch.manual_seed(77)
        g = torch.Generator().manual_seed(5)
        joint = TransducerJoint(V, E, E, J)
        pred = RNNPredictor(V, E, E, 0.1, H, L)
        ctc = CTC(V, E)
        cb = ContextBias(input_size=E, output_size=E, vocab_size=V, embedding_size=E, num_layers=1,
                         attention_heads=4, linear_units=64, num_block=2, dropout_rate=0.0,
                         bias_encoder_type="linear", bias_encoder=False, context_extractor="BLSTM",
                         num_labels=3, unified_hw_odim=E if mode == "pred" else 16)
        model = tr.Transducer(V, 0, PassThroughEncoder(), pred, joint, None, ctc, cb, ctc_weight=0.3,
                              transducer_weight=0.7, attention_weight=0.0, hw_weight=0.4, loss_mode=mode)
        model.eval()
        captured = {}
        crit = model.hw_criterion

        class Spy(torch.nn.Module):
            def forward(self, x, y, _crit=crit, _c=captured):
                _c["x"], _c["y"] = x.detach().clone(), y.detach().clone()
                return _crit(x, y)

        model.hw_criterion = Spy()
        speech = torch.randn(B, T, E, generator=g)
        speech_lengths = torch.tensor([20, 16, 20, 11])
        text = torch.randint(1, V, (B, U), generator=g)
        text_lengths = torch.tensor([7, 5, 6, 3])
        for b in range(B):
            text[b, int(text_lengths[b]):] = -1
        hw_label = torch.randint(0, 3, (B, U), generator=g)
        for b in range(B):
            hw_label[b, int(text_lengths[b]):] = -1
        context_list = torch.tensor([[0, 0, 0], [5, 9, 2], [7, 3, 0]], dtype=torch.int32)
        context_lengths = torch.tensor([1, 3, 2], dtype=torch.int32)
        out = model(speech, speech_lengths, text, text_lengths, context_list, context_lengths, hw_label)
        x = captured["x"].clone().requires_grad_(True)
        ref = crit(x, captured["y"])
        ref.backward()
        mine, gmine = orc.hw_ce_loss(captured["x"].numpy(), captured["y"].numpy())
        assert abs(mine - float(out["hw_loss"])) < 1e-6, (mode, mine, float(out["hw_loss"]))
        assert np.allclose(gmine, x.grad.numpy(), atol=1e-7), mode
        blob[f"{mode}.hw_output"] = captured["x"].numpy()
        blob[f"{mode}.labels"] = captured["y"].numpy()
        blob[f"{mode}.hw_loss"] = np.float64(float(out["hw_loss"]))
        blob[f"{mode}.grad"] = x.grad.numpy()
        blob[f"{mode}.hw_label"] = hw_label.numpy()
        blob[f"{mode}.loss"] = np.float64(float(out["loss"]))
        blob[f"{mode}.loss_rnnt"] = np.float64(float(out["loss_rnnt"]))
        blob[f"{mode}.loss_ctc"] = np.float64(float(out["loss_ctc"]))
        print(mode, "hw_output", tuple(captured["x"].shape), "labels", tuple(captured["y"].shape),
              "hw_loss", float(out["hw_loss"]), "loss", float(out["loss"]))
    blob["modes"] = np.array(["pred", "both", "sep"])
    np.savez_compressed(GOLD / "hw_loss.npz", **blob)
    print("written", GOLD / "hw_loss.npz", (GOLD / "hw_loss.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    main()
