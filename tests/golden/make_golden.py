"""make_golden.py -- regenerates tests/golden/*.npz.  Run in the BUILD container only:

    python tests/golden/make_golden.py

It needs /root/reference (for oracle/_ref, the reference's own libdsp sources compiled verbatim,
and for examples/bpsk/rrc_taps.h) and scipy (filter design).  The GPU box has neither the
reference nor a need to re-run this: the committed .npz files are the fixtures.

taps.npz     filter sets of the BASELINE.json configs
known.npz    known answers produced by the reference code on the reference's own test inputs
             (libdsp/test/test_blkconv.cxx, test_resample.py, test_decimate.py)
chain.npz    small-case outputs of the reference classes composed as a chain (cfg 4 / cfg 5 shapes)
"""
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import oracle as O  # noqa: E402


def rrc_from_reference(name):
    src = open("/root/reference/examples/bpsk/rrc_taps.h").read()
    m = re.search(r"static float %s\[\] = \{(.*?)\};" % name, src, re.S)
    return np.array([float(v.strip().rstrip("f")) for v in m.group(1).split(",") if v.strip()], np.float32)


def main():
    from scipy.signal import firwin
    O.build()
    ref = O.load("reference")
    assert ref.sfeo_kind() == b"reference"

    taps31 = np.array([4.154052476224024e-06, 6.836163811385632e-05, 0.000311697949655354, 0.0005662626936100423, -0.00019575881015043706, -0.004200732801109552, -0.013667905703186989, -0.027301859110593796, -0.03584320470690727, -0.02054252102971077, 0.04132149741053581, 0.16359668970108032, 0.33717116713523865, 0.5246731638908386, 0.670900285243988, 0.7262777090072632, 0.6709002256393433, 0.5246731042861938, 0.33717113733291626, 0.16359665989875793, 0.04132148623466492, -0.020542524755001068, -0.03584319353103638, -0.027301862835884094, -0.013667899183928967, -0.004200733732432127, -0.0001957584754563868, 0.000566263624932617, 0.0003116970765404403, 6.836453394498676e-05, 4.154052476224024e-06], np.float32)  # libdsp/test/test_resample.py:13
    taps = dict(
        lp255=firwin(255, 0.1, window="hamming").astype(np.float32),          # cfg 1 / 5 blkconv
        dec128=firwin(128, 1.0 / 8).astype(np.float32),                        # cfg 2 / 5 decimate
        rs512_l3=(3 * firwin(512, 1.0 / 625)).astype(np.float32),              # cfg 3
        rs512_l24=(24 * firwin(512, 1.0 / 625)).astype(np.float32),            # cfg 5 resample
        taps31=taps31,                                                         # reference test scripts
        rrc111=rrc_from_reference("RRC_TAPS_111"),                             # examples/bpsk/rrc_taps.h:4
        rc221=rrc_from_reference("RC_TAPS_221"),
        rrc551=rrc_from_reference("RRC_TAPS_551"),
    )
    assert len(taps["rrc111"]) == 111 and len(taps["rc221"]) == 221 and len(taps["rrc551"]) == 551
    np.savez_compressed(os.path.join(HERE, "taps.npz"), **taps)

    # ---- known answers on the reference's own test inputs -------------------------------------
    known = {}
    bc = O.Blkconv([1, 1, 1, 1, 1], 32, lib=ref)            # test_blkconv.cxx:6-34
    known["blkconv_blksize"] = np.array([bc.blksize])
    bc.buf[:28] = 1.0
    bc.process()
    known["blkconv_ones_1"] = bc.buf[:28].copy()
    bc.buf[:28] = 0.0
    bc.process()
    known["blkconv_ones_2"] = bc.buf[:28].copy()
    x0 = np.sin(0.02 * np.pi * np.arange(1024)).astype(np.float32)   # test_resample.py:16, as (float)sin(double): SURVEY.md 4.1
    for name, cls, rate in (("resample_077", O.Resample, 0.77), ("resample_177", O.Resample, 1.77),
                            ("decimate_177", O.Decimate, 1.77), ("decimate_8", O.Decimate, 8.0),
                            ("resample_8", O.Resample, 8.0)):
        o = cls(taps31, 4, 128, lib=ref)
        ys, cnt = [], []
        for b in range(8):
            n, y = o.process(x0[b * 128:(b + 1) * 128], 512, rate)
            ys.append(y[:n])
            cnt.append(n)
        known[name + "_y"] = np.concatenate(ys)
        known[name + "_counts"] = np.array(cnt)
    np.savez_compressed(os.path.join(HERE, "known.npz"), **known)

    # ---- chain goldens (small cases, reference classes composed) -------------------------------
    import fixtures as F
    chain = {}
    for name in ("cfg4_small", "cfg5_small", "cfg2_small", "cfg3_small", "cfg1_small"):
        cfg = F.CONFIGS[name]
        iq = F.make_input(cfg)
        y, n_out = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=ref)
        chain[name] = y
    np.savez_compressed(os.path.join(HERE, "chain.npz"), **chain)
    print("golden written:", {k: v.shape for k, v in chain.items()})


if __name__ == "__main__":
    main()
