"""Golden vectors for the hot-word greedy path (test infrastructure, not the product).

Runs the REAL fork greedy search (wenet/transducer/search/greedy_search.py:34-176,
context_filter_state='off') on a reference ``Transducer`` that carries a real ``ContextBias``
(wenet/transformer/context_bias.py:159-399, config as conf/encoder_bias_conformer_rnnt_*.yaml:
BLSTM extractor, linear phrase encoder, 4 heads), checks that the oracle restatement
``oracle.reference.greedy_search_hotword`` reproduces the token lists, and stores everything the
GPU test needs WITHOUT the reference tree: module weights, the phrase vectors
(forward_bias_hidden of the given and of the empty context), the biased encoder stream, hyps.
Run in the build container only: ``python oracle/gen_golden_hotword.py``.
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


def main():
    stub = types.ModuleType("turtle")
    stub.forward = None
    sys.modules.setdefault("turtle", stub)
    sys.path.insert(0, str(REF))
    import wenet.transducer.transducer as tr
    from wenet.transducer.joint import TransducerJoint
    from wenet.transducer.predictor import RNNPredictor
    from wenet.transducer.search.greedy_search import basic_greedy_search as fork_greedy
    from wenet.transformer.context_bias import ContextBias
    from wenet.transformer.ctc import CTC
    from oracle import reference as orc
    from oracle.gen_golden import PassThroughEncoder

    V, E, J, H, L = 41, 32, 48, 32, 2  # E = predictor output = context embedding size
    torch.manual_seed(20261105)
    g = torch.Generator().manual_seed(11)
    joint = TransducerJoint(V, E, E, J)
    pred = RNNPredictor(V, E, E, 0.1, H, L)
    ctc = CTC(V, E)
    cb = ContextBias(input_size=E, output_size=E, vocab_size=V, embedding_size=E, num_layers=1,
                     attention_heads=4, linear_units=64, num_block=2, dropout_rate=0.0,
                     bias_encoder_type="linear", bias_encoder=False, context_extractor="BLSTM",
                     num_labels=2, unified_hw_odim=16)
    model = tr.Transducer(V, 0, PassThroughEncoder(), pred, joint, None, ctc, cb, ctc_weight=0.3,
                          transducer_weight=0.7, attention_weight=0.0, hw_weight=0.0, loss_mode="pred")
    model.eval()
    with torch.no_grad():  # spread the logits so that symbols are emitted and argmax margins are wide
        joint.ffn_out.weight.mul_(4.0)
        joint.ffn_out.bias[0] += 1.5
        pred.embed.weight.mul_(3.0)
        for p in cb.predictor_bias.parameters():
            p.mul_(2.0)
    # phrase list: first entry is the 1-token "no phrase" entry the fork's lists start with
    context_list = torch.tensor([[0, 0, 0], [5, 9, 2], [7, 3, 0], [11, 0, 0]], dtype=torch.int32)
    context_lengths = torch.tensor([1, 3, 2, 1], dtype=torch.int32)
    Bs, Ts = 5, 30
    encs = torch.randn(Bs, Ts, E, generator=g) * 1.5
    lens = torch.tensor([30, 22, 30, 17, 26])
    n_steps = 4
    with torch.no_grad():
        bias_hidden = cb.forward_bias_hidden(context_list, context_lengths)
        empty = torch.zeros((1, 1), dtype=torch.int)
        bias_hidden_empty = cb.forward_bias_hidden(empty, context_lengths[0].unsqueeze(0))
        enc_biased, _ = cb.forward_encoder_bias(bias_hidden, encs)
    hyps_all = []
    for b in range(Bs):
        e1 = encs[b:b + 1, :int(lens[b])]
        with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
            out = fork_greedy(model, e1, lens[b], context_list, context_lengths, n_steps=n_steps,
                              context_filter_state="off",
                              context_decoder_labels_padded=torch.IntTensor([[0]]))
        want = out[0][0]
        mine = orc.greedy_search_hotword(model, enc_biased[b:b + 1, :int(lens[b])], int(lens[b]),
                                         bias_hidden_empty, n_steps=n_steps)
        assert want == mine, (b, want, mine)
        hyps_all.append(want)
        print(f"utt {b}: T={int(lens[b])} -> {len(want)} tokens {want[:10]}")
    # loss_mode 'both': basic_greedy_search_both (greedy_search.py:297-430), filter 'off': the predictor
    # bias attends to the GIVEN phrase list (:357)
    from wenet.transducer.search.greedy_search import basic_greedy_search_both as fork_greedy_both
    hyps_both = []
    for b in range(Bs):
        e1 = encs[b:b + 1, :int(lens[b])]
        with torch.no_grad(), contextlib.redirect_stdout(io.StringIO()):
            out = fork_greedy_both(model, e1, lens[b], context_list, context_lengths, n_steps=n_steps,
                                   context_filter_state="off",
                                   context_decoder_labels_padded=torch.IntTensor([[0]]))
        want = out[0][0]
        mine = orc.greedy_search_hotword(model, enc_biased[b:b + 1, :int(lens[b])], int(lens[b]),
                                         bias_hidden, n_steps=n_steps)
        assert want == mine, ("both", b, want, mine)
        hyps_both.append(want)
    print("both-mode hyps differ from pred-mode:", hyps_both != hyps_all)
    assert any(hyps_all), "degenerate fixture: nothing emitted"
    # the same model WITHOUT predictor bias must decode differently somewhere (the path matters)
    plain = [orc.greedy_search(model, enc_biased[b:b + 1, :int(lens[b])], int(lens[b]), n_steps=n_steps)
             for b in range(Bs)]
    assert plain != hyps_all, "predictor bias had no effect: fixture does not exercise the path"
    width = max(1, max(len(h) for h in hyps_all))
    arr = np.full((Bs, width), -1, dtype=np.int64)
    for b, h in enumerate(hyps_all):
        arr[b, :len(h)] = h
    wb = max(1, max(len(h) for h in hyps_both))
    arr_both = np.full((Bs, wb), -1, dtype=np.int64)
    for b, h in enumerate(hyps_both):
        arr_both[b, :len(h)] = h
    blob = dict(hyps_both=arr_both, dims=np.array([V, E, E, J, H, L]), enc=encs.numpy(), enc_biased=enc_biased.numpy(),
                lens=lens.numpy(), hyps=arr, n_steps=np.int64(n_steps), heads=np.int64(4),
                bias_hidden=bias_hidden.numpy(), bias_hidden_empty=bias_hidden_empty.numpy(),
                context_list=context_list.numpy(), context_lengths=context_lengths.numpy())
    for prefix, mod in (("w.predictor.", pred), ("w.joint.", joint), ("w.ctc.", ctc), ("cb.", cb)):
        for k, v in mod.state_dict().items():
            if prefix == "cb." and not k.startswith(("predictor_bias", "encoder_bias_combine",
                                                     "encdoer_bias", "encoder_bias.")):
                continue
            blob[prefix + k] = v.detach().numpy()
    GOLD.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(GOLD / "hotword_greedy.npz", **blob)
    print("written", GOLD / "hotword_greedy.npz", (GOLD / "hotword_greedy.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    os.environ.setdefault("PYTHONDONTWRITEBYTECODE", "1")
    main()
