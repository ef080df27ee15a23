"""wenet_celoss_b200 — B200-native (sm_100a) transducer hot path for WeNet-style models.

Public surface mirrors the reference call sites (Vindicator645/wenet-celoss):
  rnnt_loss(logits, targets, logit_lengths, target_lengths, blank, clamp, reduction)
      torchaudio.functional.rnnt_loss as used at wenet/transducer/transducer.py:142
  CTC(odim, encoder_output_size).forward(hs_pad, hlens, ys_pad, ys_lens)
      wenet/transformer/ctc.py:46
  fused_joint_rnnt_loss(joint, enc_out, pred_out, targets, logit_lengths, target_lengths, blank, reduction)
      wenet/transducer/transducer.py:132-147 (joint + rnnt_loss without materialised logits)
  greedy_search_batch / basic_greedy_search, prefix_beam_search_batch / PrefixBeamSearch
      wenet/transducer/search/greedy_search.py:34, prefix_beam_search.py:42
  ctc_prefix_beam_search_batch(ctc_probs, lens, beam_size)
      wenet/transformer/asr_model.py:326-409 (CTC prefix beam search, batched on the device)
  patch_transducer(model)  rebinds Transducer.forward / greedy_search / beam_search
Everything is backed by libb200_transducer.so (C ABI in include/b200_transducer.h);
there is no CPU or PyTorch fallback.
"""
from ._lib import B200Error, launch_count, lib  # noqa: F401
from .ctc import CTC, ctc_loss, patch_ctc  # noqa: F401
from .ctc_search import ctc_prefix_beam_search_batch  # noqa: F401
from .hw_loss import hw_ce_loss  # noqa: F401
from .joint import fused_joint_rnnt_loss, joint_logprobs  # noqa: F401
from .search import (PrefixBeamSearch, Sequence, basic_greedy_search,  # noqa: F401
                     greedy_search_batch, prefix_beam_search_batch)
from .transducer import add_blank, patch_transducer, rnnt_targets  # noqa: F401
from .rnnt import rnnt_grad_coef, rnnt_lattice, rnnt_loss  # noqa: F401

__all__ = ["patch_transducer", "add_blank", "rnnt_targets", "prefix_beam_search_batch", "PrefixBeamSearch", "Sequence", "greedy_search_batch", "basic_greedy_search", "fused_joint_rnnt_loss", "CTC", "ctc_loss", "patch_ctc", "hw_ce_loss", "ctc_prefix_beam_search_batch", "rnnt_loss", "rnnt_lattice", "rnnt_grad_coef", "B200Error", "lib", "launch_count"]
