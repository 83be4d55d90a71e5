"""Streaming front door: sfe_callback-compatible producer -> pinned ring -> GPU chain,
against the reference classes fed the same accepted bytes.  Behaviour mirrored from
gr-simplefe/lib/source_c_impl.cc:84-108 (write_data) and :134-153 (work)."""
import numpy as np
import pytest

import fixtures as F


@pytest.mark.gpu
def test_rx_feed_streams_packets_through_the_chain(gpu_lib, oracle_lib, capfd):
    from oracle import oracle as O
    from simplefe_b200.chain import RxFeed
    cfg = dict(F.CONFIGS["cfg5_small"], in_fmt=F.IN_U8, S=1, N=8 * 4096)
    iq = F.make_input(cfg)[0]                      # (N, 2) uint8: the bytes the board sends, I then Q
    raw = iq.tobytes()
    block = 4096
    feed = RxFeed(ring_bytes=3 * 2 * block, block_samples=block, blkconv=cfg["blkconv"], stage_a=cfg["stage_a"], stage_b=cfg["stage_b"])
    got = []
    pkt = 1024                                      # ISO packet size
    accepted = bytearray()
    pos = 0
    k = 0
    while pos < len(raw):
        p = raw[pos:pos + pkt]
        assert feed.callback(p) == 0
        accepted += p
        pos += pkt
        k += 1
        if k == 5:
            assert feed.callback(b"\x01\x02\x03") == 0          # odd length: discarded with the reference's message
        if k % 7 == 0:                                           # the consumer runs now and then
            got.append(feed.pump(cap=100000))
    got.append(feed.pump(cap=100000, wait_ms=10))
    y = np.concatenate(got)
    assert feed.stat(1) == 1 and feed.stat(0) == 0 and feed.stat(2) == len(accepted) and feed.stat(3) == 0
    assert "odd number" in capfd.readouterr().out
    x = np.frombuffer(bytes(accepted), np.uint8).reshape(1, -1, 2)
    ref, _ = O.chain_run(x, F.IN_U8, cfg["blkconv"], cfg["stage_a"], cfg["stage_b"], lib=oracle_lib)
    assert y.shape == ref[0].shape
    nrms, peak = F.err_metrics(y, ref[0])
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.gpu
def test_rx_feed_overflow_prints_O_and_drops(gpu_lib, capfd):
    from simplefe_b200.chain import RxFeed
    cfg = F.CONFIGS["cfg2_small"]
    feed = RxFeed(ring_bytes=4096, block_samples=2048, stage_a=cfg["stage_a"])
    assert feed.callback(bytes(4096)) == 0
    assert feed.callback(bytes(2)) == 0             # ring full: dropped, like source_c_impl.cc:98-100
    assert feed.stat(0) == 1 and feed.stat(2) == 4096
    assert "O" in capfd.readouterr().err
    assert len(feed.pump(cap=4096)) > 0 and feed.stat(3) == 0
    assert feed.callback(bytes(2)) == 0 and feed.stat(0) == 1
