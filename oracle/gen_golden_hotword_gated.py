"""Golden vectors for the fork's DEFAULT greedy path: hot-word gate with "go-back"
(test infrastructure, not the product).

Runs the REAL ``basic_greedy_search`` (wenet/transducer/search/greedy_search.py:34-176) with
``context_filter_state='on'`` — the default of ``Transducer.greedy_search`` (transducer.py:525)
and of ``wenet/bin/recognize.py:139-141`` — on a reference ``Transducer`` with a real
``ContextBias`` whose ``unified_hw_odim`` equals the embedding size (as
conf/encoder_bias_conformer_rnnt_4_head_bi_1_layer_pred.yaml:40,51; ``forward_hw_pred`` needs it).
Checks that ``oracle.reference.greedy_search_hotword_gated`` reproduces token lists, gate
decisions and the returned edit distance, requires that the fixture really goes back, and
stores what the GPU test needs without the reference tree.
Run in the build container only: ``python oracle/gen_golden_hotword_gated.py``.
"""
from __future__ import annotations

import contextlib
import io
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


def make(mode: str):
    stub = types.ModuleType("turtle")
    stub.forward = None
    sys.modules.setdefault("turtle", stub)
    sys.path.insert(0, str(REF))
    import wenet.transducer.transducer as tr
    from wenet.transducer.joint import TransducerJoint
    from wenet.transducer.predictor import RNNPredictor
    from wenet.transducer.search import greedy_search as gs_mod
    from wenet.transformer.context_bias import ContextBias
    from wenet.transformer.ctc import CTC
    from oracle import reference as orc
    from oracle.gen_golden import PassThroughEncoder

    V, E, J, H, L = 41, 32, 48, 32, 2
    torch.manual_seed(20261107)
    g = torch.Generator().manual_seed(13)
    joint = TransducerJoint(V, E, E, J)
    pred = RNNPredictor(V, E, E, 0.1, H, L)
    ctc = CTC(V, E)
    cb = ContextBias(input_size=E, output_size=E, vocab_size=V, embedding_size=E, num_layers=1,
                     attention_heads=4, linear_units=64, num_block=2, dropout_rate=0.0,
                     bias_encoder_type="linear", bias_encoder=False, context_extractor="BLSTM",
                     num_labels=2, unified_hw_odim=E if mode == "pred" else 16, unified_hw_heads=4)
    model = tr.Transducer(V, 0, PassThroughEncoder(), pred, joint, None, ctc, cb, ctc_weight=0.3,
                          transducer_weight=0.7, attention_weight=0.0, hw_weight=0.0, loss_mode=mode)
    model.eval()
    with torch.no_grad():  # wide argmax margins, frequent emissions, a gate that flips
        joint.ffn_out.weight.mul_(4.0)
        joint.ffn_out.bias[0] += 1.5
        pred.embed.weight.mul_(3.0)
        for p in cb.predictor_bias.parameters():
            p.mul_(2.0)
        for p in cb.hw_bias.parameters():
            p.mul_(3.0)
        cb.hw_output_layer.weight.mul_(60.0 if mode == "pred" else 4.0)
    context_list = torch.tensor([[0, 0, 0], [5, 9, 2], [7, 3, 0], [11, 0, 0], [4, 8, 15]], dtype=torch.int32)
    context_lengths = torch.tensor([1, 3, 2, 1, 3], dtype=torch.int32)
    Bs, Ts = 6, 40
    encs = torch.randn(Bs, Ts, E, generator=g) * 1.5
    lens = torch.tensor([40, 22, 40, 17, 31, 36])
    n_steps = 3
    labels = torch.randint(0, 2, (1, 12), generator=g).int()   # context_decoder_labels_padded
    with torch.no_grad():
        bias_hidden = cb.forward_bias_hidden(context_list, context_lengths)
        empty = torch.zeros((1, 1), dtype=torch.int)
        bias_hidden_empty = cb.forward_bias_hidden(empty, context_lengths[0].unsqueeze(0))
        enc_biased, enc_bias = cb.forward_encoder_bias(bias_hidden, encs)
        enc_empty, _ = cb.forward_encoder_bias(bias_hidden_empty, encs)

    # centre the gate: record hw outputs over one reference pass, shift the label-1 bias by the median
    fork = gs_mod.basic_greedy_search if mode == "pred" else gs_mod.basic_greedy_search_both
    hw_name = "forward_hw_pred" if mode == "pred" else "forward_hw_pred_both"
    seen = []
    real_hw = getattr(cb, hw_name)

    def recording(*a, **k):
        o = real_hw(*a, **k)
        seen.append(o.detach().reshape(-1).clone())
        return o
    setattr(cb, hw_name, recording)
    for b in range(Bs):
        with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
            fork(model, encs[b:b + 1, :int(lens[b])], lens[b], context_list,
                                       context_lengths, n_steps=n_steps, context_filter_state="on",
                                       context_decoder_labels_padded=labels)
    d = torch.stack(seen)
    mg = torch.unique((d[:, 1] - d[:, 0]).round(decimals=4))   # sorted; one entry per distinct frame
    with torch.no_grad():   # midway between the two middle samples: no sample sits on the boundary
        cb.hw_output_layer.bias[1] -= float(mg[len(mg) // 2 - 1] + mg[len(mg) // 2]) / 2
    print("gate margin spread:", float(mg.std()))
    seen.clear()            # keeps recording: the margins of the final pass are checked below

    hyps_all, res_all, dist_all, n_back = [], [], [], 0
    for b in range(Bs):
        n = int(lens[b])
        e1 = encs[b:b + 1, :n]
        with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
            out = fork(model, e1, lens[b], context_list, context_lengths, n_steps=n_steps,
                       context_filter_state="on", context_decoder_labels_padded=labels)
        want_hyp, want_dist = out[0][0], float(out[1])
        eb = None if mode == "pred" else enc_bias[b:b + 1, :n]
        hyp, res = orc.greedy_search_hotword_gated(model, enc_biased[b:b + 1, :n], enc_empty[b:b + 1, :n], n,
                                                   bias_hidden, bias_hidden_empty, n_steps=n_steps, enc_bias=eb)
        assert want_hyp == hyp, (b, want_hyp, hyp)
        if mode == "pred":   # basic_greedy_search_both does not return its decisions, only their distance
            assert list(out[2]) == res, (b, out[2], res)
        assert want_dist == orc.edit_distance(labels[0].tolist(), res)
        hyps_all.append(hyp)
        res_all.append(res)
        dist_all.append(want_dist)
        print(f"utt {b}: T={n} -> {len(hyp)} tokens, decisions {res}")
    setattr(cb, hw_name, real_hw)
    d = torch.stack(seen)
    min_margin = float((d[:, 1] - d[:, 0]).abs().min())
    print("smallest gate margin in the fixture:", min_margin)
    assert min_margin > 1e-4, "a gate decision sits within rounding of the boundary: change the seed"
    # the 'off' decode of the same model must differ (the gate matters) and both decisions occur
    off = [orc.greedy_search_hotword(model, enc_biased[b:b + 1, :int(lens[b])], int(lens[b]),
                                     bias_hidden_empty if mode == "pred" else bias_hidden, n_steps=n_steps)
           for b in range(Bs)]
    assert off != hyps_all, "gate had no effect: fixture does not exercise the path"
    flat = [r for rs in res_all for r in rs]
    assert 0 in flat and 1 in flat
    # count go-backs through the predictor calls of the (verified) restatement
    import oracle.reference as R
    for b in range(Bs):
        n = int(lens[b])
        cnt = {"calls": 0}
        real = model.predictor.forward_step

        def counting(*a, _real=real, **k):
            cnt["calls"] += 1
            return _real(*a, **k)
        model.predictor.forward_step = counting
        try:
            R.greedy_search_hotword_gated(model, enc_biased[b:b + 1, :n], enc_empty[b:b + 1, :n], n,
                                          bias_hidden, bias_hidden_empty, n_steps=n_steps,
                                          enc_bias=None if mode == "pred" else enc_bias[b:b + 1, :n])
        finally:
            model.predictor.forward_step = real
        # every go-back costs one extra call and removes one decision: calls = len(res) + 2 * go_backs
        n_back += (cnt["calls"] - len(res_all[b])) // 2
    print("go-backs in the fixture:", n_back)
    assert n_back >= 3, "fixture does not exercise the go-back"

    def pack(lists):
        w = max(1, max(len(h) for h in lists))
        arr = np.full((len(lists), w), -1, dtype=np.int64)
        for i, h in enumerate(lists):
            arr[i, :len(h)] = h
        return arr
    blob = dict(dims=np.array([V, E, E, J, H, L]), enc=encs.numpy(), enc_biased=enc_biased.numpy(),
                enc_empty=enc_empty.numpy(), lens=lens.numpy(), hyps=pack(hyps_all), decisions=pack(res_all),
                dist=np.array(dist_all), labels=labels.numpy(), n_steps=np.int64(n_steps), heads=np.int64(4),
                go_backs=np.int64(n_back), bias_hidden=bias_hidden.numpy(),
                bias_hidden_empty=bias_hidden_empty.numpy(), context_list=context_list.numpy(),
                context_lengths=context_lengths.numpy(), hyps_off=pack(off), enc_bias=enc_bias.numpy())
    for prefix, mod in (("w.predictor.", pred), ("w.joint.", joint), ("w.ctc.", ctc), ("cb.", cb)):
        for k, v in mod.state_dict().items():
            blob[prefix + k] = v.detach().numpy()
    GOLD.mkdir(parents=True, exist_ok=True)
    out = GOLD / ("hotword_greedy_gated.npz" if mode == "pred" else "hotword_greedy_gated_both.npz")
    np.savez_compressed(out, **blob)
    print("written", out, out.stat().st_size, "bytes")


def main():
    make("pred")
    make("both")


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    main()
