"""TEST INFRASTRUCTURE (not the product): golden vectors for n-best rescoring.

Runs the REAL ``Transducer.transducer_attention_rescoring`` of /root/reference
(wenet/transducer/transducer.py:378-513) - real RNNPredictor / TransducerJoint / CTC and a real
(Bi)TransformerDecoder - on seeded inputs, records what its rescoring step received and returned
(n-best lists, search scores, ``_cal_transducer_score``, the decoder logits of the n-best, the best
hypothesis and its score), asserts that the restatement ``oracle.reference.attention_rescoring`` gives
the same answer, and writes tests/golden/rescore.npz.

    python oracle/gen_golden_rescore.py        (needs /root/reference; run in the build container)
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.dont_write_bytecode = True
from oracle.gen_golden import GOLD, PassThroughEncoder, import_reference, make_identity_bias, sd_np  # noqa: E402


def main():
    tr = import_reference()
    from wenet.transducer.joint import TransducerJoint
    from wenet.transducer.predictor import RNNPredictor
    from wenet.transformer.ctc import CTC
    from wenet.transformer.decoder import BiTransformerDecoder

    from oracle import reference as orc

    V, E, P, J, H, L = 41, 16, 16, 64, 16, 2  # join dim: a multiple of 64 (the fused joint kernels)
    torch.manual_seed(20261105)
    g = torch.Generator().manual_seed(11)
    pred = RNNPredictor(V, E, P, 0.1, H, L)
    joint = TransducerJoint(V, E, P, J)
    ctc = CTC(V, E)
    dec = BiTransformerDecoder(V, E, attention_heads=2, linear_units=32, num_blocks=1, r_num_blocks=1,
                               dropout_rate=0.0, positional_dropout_rate=0.0)
    model = tr.Transducer(V, 0, PassThroughEncoder(), pred, joint, dec, ctc, make_identity_bias(),
                          ctc_weight=0.3, transducer_weight=0.5, attention_weight=0.2, hw_weight=0.0,
                          loss_mode='pred', reverse_weight=0.3)
    model.eval()
    with torch.no_grad():  # spread the logits so that beams hold distinct, non-trivial hypotheses
        joint.ffn_out.weight.mul_(3.0)
        ctc.ctc_lo.weight.mul_(3.0)
        pred.embed.weight.mul_(3.0)

    om = orc.OracleModel(V, E, P, J, E, H, L, blank=0)
    om.predictor.load_state_dict(pred.state_dict())
    om.joint.load_state_dict(joint.state_dict())
    om.ctc.load_state_dict(ctc.state_dict())
    om.eval()

    out = {"dims": np.array([V, E, P, J, H, L], dtype=np.int64), "sos_eos": np.array([model.sos, model.eos])}
    for k, v in {**sd_np(pred, "predictor."), **sd_np(joint, "joint."), **sd_np(ctc, "ctc.")}.items():
        out["w." + k] = v

    cases = [
        dict(T=14, beam=4, kind="transducer", rw=0.0, cw=0.2, aw=0.5, tw=0.3, scw=0.3, stw=0.7),
        dict(T=17, beam=5, kind="transducer", rw=0.3, cw=0.1, aw=0.6, tw=0.3, scw=0.3, stw=0.7),
        dict(T=12, beam=3, kind="ctc", rw=0.0, cw=0.3, aw=0.4, tw=0.3, scw=1.0, stw=0.0),
        dict(T=15, beam=4, kind="ctc", rw=0.4, cw=0.5, aw=0.5, tw=0.0, scw=1.0, stw=0.0),
    ]
    for ci, c in enumerate(cases):
        speech = torch.randn(1, c["T"], E, generator=g) * 2.0
        lens = torch.tensor([c["T"]])
        rec = {}
        td_orig, attn_orig = model._cal_transducer_score, model._cal_attn_score

        def td_hook(encoder_out, encoder_mask, hyps_lens, hyps_pad):
            r = td_orig(encoder_out, encoder_mask, hyps_lens, hyps_pad)
            rec.update(td=r.detach().clone(), hyps_pad=hyps_pad.clone(), hyps_lens=hyps_lens.clone(),
                       enc_rep=encoder_out.clone())
            return r

        def attn_hook(encoder_out, encoder_mask, hyps_pad, hyps_lens):
            # the decoder's raw logits for the n-best (what a stand-in decoder replays in the GPU test)
            from wenet.utils.common import add_sos_eos, reverse_pad_list
            hp, _ = add_sos_eos(hyps_pad, model.sos, model.eos, model.ignore_id)
            rhp = reverse_pad_list(hyps_pad, hyps_lens, model.ignore_id)
            rhp, _ = add_sos_eos(rhp, model.sos, model.eos, model.ignore_id)
            d, rd, _ = model.decoder(encoder_out, encoder_mask, hp, hyps_lens + 1, rhp, model.reverse_weight)
            rec.update(dec=d.detach().clone(), rdec=rd.detach().clone() if torch.is_tensor(rd) else None,
                       ys_in=hp.clone(), r_ys_in=rhp.clone())
            return attn_orig(encoder_out, encoder_mask, hyps_pad, hyps_lens)

        model._cal_transducer_score, model._cal_attn_score = td_hook, attn_hook
        model.reverse_weight = c["rw"]
        with torch.no_grad():
            best_hyp, best_score = model.transducer_attention_rescoring(
                speech, lens, c["beam"], reverse_weight=c["rw"], ctc_weight=c["cw"], attn_weight=c["aw"],
                transducer_weight=c["tw"], search_ctc_weight=c["scw"], search_transducer_weight=c["stw"],
                beam_search_type=c["kind"])
        model._cal_transducer_score, model._cal_attn_score = td_orig, attn_orig
        hyps = [[int(x) for x in row[:n]] for row, n in zip(rec["hyps_pad"].tolist(), rec["hyps_lens"].tolist())]
        # the search scores the reference combined (re-run the search the same way)
        with torch.no_grad():
            if c["kind"] == "transducer":
                model.bs = None
                model.init_bs()
                beam, _ = model.bs.prefix_beam_search(speech, lens, beam_size=c["beam"], ctc_weight=c["scw"],
                                                      transducer_weight=c["stw"])
                beam_score = [float(s.score) for s in beam]
                assert [s.hyp[1:] for s in beam] == hyps
                want = orc.prefix_beam_search(om, speech, beam_size=c["beam"], ctc_weight=c["scw"],
                                              transducer_weight=c["stw"])
                assert [s.hyp[1:] for s in want] == hyps, "oracle beam differs from the reference"
            else:
                nbest, _ = model._ctc_prefix_beam_search(speech, lens, beam_size=c["beam"])
                beam_score = [float(h[1]) for h in nbest]
                assert [list(h[0]) for h in nbest] == hyps
        td = rec["td"]
        td_or = orc.transducer_score(om, rec["enc_rep"], torch.full((c["beam"],), c["T"]), rec["hyps_pad"],
                                     rec["hyps_lens"])
        assert torch.allclose(td, td_or, rtol=1e-5, atol=1e-5), (td, td_or)
        dlp = torch.log_softmax(rec["dec"], -1)
        rlp = torch.log_softmax(rec["rdec"], -1) if rec["rdec"] is not None and rec["rdec"].dim() == 3 else None
        bi, bs, attn = orc.attention_rescoring(hyps, beam_score, td.tolist(), dlp.numpy(),
                                               None if rlp is None else rlp.numpy(), model.eos, c["rw"], c["aw"],
                                               c["cw"], c["tw"])
        assert hyps[bi] == list(best_hyp), (hyps[bi], best_hyp)
        assert abs(bs - float(best_score)) <= 1e-5 * max(1.0, abs(bs)), (bs, float(best_score))
        assert len({tuple(h) for h in hyps}) == len(hyps)
        pre = f"c{ci}."
        out[pre + "speech"] = speech.numpy()
        out[pre + "params"] = np.array([c["beam"], c["rw"], c["cw"], c["aw"], c["tw"], c["scw"], c["stw"]])
        out[pre + "kind"] = np.array(c["kind"])
        out[pre + "hyps_pad"] = rec["hyps_pad"].numpy()
        out[pre + "hyps_lens"] = rec["hyps_lens"].numpy()
        out[pre + "beam_score"] = np.array(beam_score)
        out[pre + "td_score"] = td.numpy()
        out[pre + "dec"] = rec["dec"].numpy()
        if rlp is not None:
            out[pre + "rdec"] = rec["rdec"].numpy()
        out[pre + "ys_in"] = rec["ys_in"].numpy()
        out[pre + "r_ys_in"] = rec["r_ys_in"].numpy()
        out[pre + "attn"] = np.array(attn)
        out[pre + "best_hyp"] = np.array(list(best_hyp), dtype=np.int64)
        out[pre + "best_score"] = np.array(float(best_score))
        out[pre + "best_index"] = np.array(bi)
        print(f"case {ci}: {c['kind']} beam {c['beam']} best #{bi} {list(best_hyp)} score {float(best_score):.5f}")
    out["ncases"] = np.array(len(cases))
    np.savez_compressed(GOLD / "rescore.npz", **out)
    print("wrote", GOLD / "rescore.npz")


if __name__ == "__main__":
    main()
