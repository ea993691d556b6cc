"""Regenerates tests/golden/*.npz from the CPU oracle (oracle/r2p1d_ref.py).

The reference ships no golden vectors for this path (no tests, no checkpoint, missing submodule),
so these fixtures pin the ORACLE, not the reference: they catch accidental changes to the
restatement and give the GPU parity tests fixed targets.  Run from the repo root:

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import r2p1d_ref as R  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def tap_stats(taps):
    names = sorted(taps)
    return names, np.array([[float(taps[n].double().abs().mean()), float(taps[n].double().sum()),
                             float(taps[n].abs().max())] for n in names], dtype=np.float64)


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    clips = R.synthetic_clips(2, seed=1234)
    out = {}
    for tag, harden in (('default', False), ('hardened', True)):
        model = R.build(seed=0, harden=harden)
        taps = {}
        logits = R.forward_u8(model, clips, taps)
        names, stats = tap_stats(taps)
        out['r18_%s_logits' % tag] = logits.numpy().astype(np.float32)
        out['r18_%s_tap_names' % tag] = np.array(names)
        out['r18_%s_tap_stats' % tag] = stats
    # the reference's own warm-up input distribution (models/r2p1d/model.py:68)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(1, 3, 8, 112, 112, generator=g)
    out['r18_default_randn_logits'] = R.forward(R.build(seed=0), x).numpy().astype(np.float32)
    # R(2+1)D-34 in the runner's 5-entry layer_sizes form
    m34 = R.build(layer_sizes=[0, 3, 4, 6, 3], seed=0, harden=True)
    out['r34_hardened_logits'] = R.forward_u8(m34, clips[:1]).numpy().astype(np.float32)
    # a layer range that neither starts at the stem nor ends at the classifier
    g = torch.Generator().manual_seed(11)
    x34 = torch.randn(2, 64, 8, 56, 56, generator=g).abs()
    m_mid = R.build(3, 4, seed=0, harden=True)
    y = R.forward(m_mid, x34)
    out['r18_hardened_layers3to4_stats'] = np.array(
        [float(y.double().abs().mean()), float(y.double().sum()), float(y.abs().max())])
    np.savez_compressed(os.path.join(OUT, 'r2p1d_oracle.npz'), **out)
    for k, v in out.items():
        print(k, v.shape, v.dtype)


if __name__ == '__main__':
    main()
