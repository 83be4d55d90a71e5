"""GNU Radio source block over the GPU chain: gr::simplefe::gpu_source_c, compiled against
stand-in gnuradio/boost headers (tests/cpp/stubs; GNU Radio itself is absent from this image), driven like the
scheduler drives gr-simplefe/lib/source_c_impl.cc -- a producer thread calling rx_callback with ISO packets, work()
asked for fixed-size item counts -- and compared with the reference classes fed the same bytes."""
import os
import struct
import subprocess

import numpy as np
import pytest

import fixtures as F

HERE = os.path.dirname(os.path.abspath(__file__))
EXE = os.path.join(HERE, "cpp", "_bin", "gr_block_check")


def _stage_bytes(st):
    if st is None:
        return struct.pack("<iiiif", 0, 0, 1, 0, 1.0)
    t = np.ascontiguousarray(st["taps"], np.float32)
    return struct.pack("<iiiif", int(st["klass"]), len(t), int(st["up"]), int(st["blk"]), float(st["rate"])) + t.tobytes()


def _chain_bytes(cfg):
    bc = cfg.get("blkconv")
    if bc is None:
        head = struct.pack("<ii", 0, 0)
    else:
        t = np.ascontiguousarray(bc[0], np.float32)
        head = struct.pack("<ii", len(t), int(bc[1])) + t.tobytes()
    return head + _stage_bytes(cfg.get("stage_a")) + _stage_bytes(cfg.get("stage_b"))


def _reference_fed_per_block(O, lib, iq, cfg, block):
    """the reference classes driven the way the block's feed drives the chain: one chain call per `block` device samples,
    every decimate/resample object getting that call's samples in one process() (with a fractional rate the float32
    time walk of decimate.cxx:68-104 depends on where the calls are cut, so the reference must be cut the same way)"""
    comps = []
    for c in range(2):
        f = O.u8_to_f32(iq[:, c], lib=lib)
        if cfg.get("blkconv") is not None:
            f = O.Blkconv(cfg["blkconv"][0], cfg["blkconv"][1], lib=lib).filter(f)
        objs = []
        for st in (cfg.get("stage_a"), cfg.get("stage_b")):
            if st is not None:
                cls = O.Decimate if st["klass"] == F.KLASS_DECIMATE else O.Resample
                objs.append((cls(st["taps"], st["up"], st["blk"], lib=lib), st))
        outs = []
        for off in range(0, len(f), block):
            v = f[off:off + block]
            for obj, st in objs:
                assert len(v) <= st["blk"]
                n, y = obj.process(v, int(np.floor(np.float32(st["blk"]) / np.float32(st["rate"]))) + 2, st["rate"])
                v = y[:n]
            outs.append(v)
        comps.append(np.concatenate(outs))
    return (comps[0] + 1j * comps[1]).astype(np.complex64)


def test_gr_block_sources_exist_and_mirror_the_reference_block():
    hdr = open(os.path.join(HERE, "..", "include", "simplefe", "gpu_source_c.h")).read()
    for needle in ("virtual public gr::sync_block", "boost::shared_ptr<gpu_source_c> sptr", "static sptr make(", "rx_callback("):
        assert needle in hdr
    impl = open(os.path.join(HERE, "..", "simplefe_b200", "gr", "gpu_source_c_impl.cc")).read()
    assert "oracle" not in impl and "sfe_rx_feed_pump" in impl
    # the other three block shapes of gr-simplefe (sink_c, sink_f, source_f) and their GRC descriptions
    snk = open(os.path.join(HERE, "..", "include", "simplefe", "gpu_sink.h")).read()
    for needle in ("class gpu_sink_c : virtual public gr::sync_block", "class gpu_sink_f : virtual public gr::sync_block",
                   "boost::shared_ptr<gpu_sink_c> sptr", "boost::shared_ptr<gpu_sink_f> sptr", "tx_callback("):
        assert needle in snk
    srcf = open(os.path.join(HERE, "..", "include", "simplefe", "gpu_source_f.h")).read()
    assert "class gpu_source_f : virtual public gr::sync_block" in srcf and "int channel" in srcf
    for b in ("gpu_source_c", "gpu_source_f", "gpu_sink_c", "gpu_sink_f"):
        xml = open(os.path.join(HERE, "..", "simplefe_b200", "gr", "grc", f"simplefe_{b}.xml")).read()
        assert f"<key>simplefe_{b}</key>" in xml and f"simplefe.{b}(" in xml
    for f in ("gpu_sink_impl.cc", "gpu_source_f_impl.cc"):
        assert "oracle" not in open(os.path.join(HERE, "..", "simplefe_b200", "gr", f)).read()


SINK_EXE = os.path.join(HERE, "cpp", "_bin", "gr_sink_check")


@pytest.mark.gpu
@pytest.mark.parametrize("mode,pkt", [("sink_c", 500), ("sink_f", 1280), ("sink_f", 505)])
def test_gr_sink_blocks_put_the_reference_bytes_on_the_wire(gpu_lib, oracle_lib, tmp_path, mode, pkt):
    """work() in ragged item counts on one thread, tx_callback packets on another: the wire bytes are those of the
    reference's fill_tx_buffer (sink_c_impl.cc:118-144 = bpsk.cxx:76-101) applied to the same samples, and -- where the
    reference itself is compiled in -- those its own ring_buffer<float> + tx_callback (bpsk.cxx:104-119) deliver"""
    from oracle import oracle as O
    if not os.path.exists(SINK_EXE):
        subprocess.run(["python", os.path.join(HERE, "cpp", "build_cpp.py")], check=True)
    rng = np.random.default_rng(21)
    n = 50003 if mode == "sink_f" else 2 * 25001
    x = np.clip(rng.normal(0, 0.4, n), -1.0, 1.0).astype(np.float32)
    x[:8] = [1.0, -1.0, 0.0, 0.5, -0.5, 0.999, -0.999, 1e-3]
    (tmp_path / "in.f32").write_bytes(x.tobytes())
    r = subprocess.run([SINK_EXE, mode, str(tmp_path / "in.f32"), str(tmp_path / "out.bytes"), str(pkt), "-"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "underflows 1" in r.stdout and r.stderr.count("U") == 1     # only the deliberate one
    wire = np.fromfile(tmp_path / "out.bytes", np.uint8)
    whole = len(x) // 4 * 4
    want = O.f32_to_dac10(x[:whole], lib=oracle_lib)
    n_pk = len(want) // pkt
    assert len(wire) == n_pk * (pkt // 5 * 5) and n_pk > 10
    assert np.array_equal(wire, want[: len(wire)])
    ref = O.ref_tx_packets(x[:whole], whole + 8, pkt, lib=oracle_lib)
    if ref is not None and pkt % 5 == 0:
        assert np.array_equal(wire, ref[: len(wire)]), "same bytes as the reference's ring_buffer<float> + tx_callback"


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cfg2_small", "cfg5_small"])
def test_gr_source_f_one_channel(gpu_lib, oracle_lib, tmp_path, name):
    """one real channel (source_f_impl.cc:120-129: one byte per sample) through the chain: equals the reference classes
    fed that channel's floats (the imaginary part of the stream is the format's zero and stays zero)"""
    from oracle import oracle as O
    if not os.path.exists(SINK_EXE):
        subprocess.run(["python", os.path.join(HERE, "cpp", "build_cpp.py")], check=True)
    cfg = dict(F.CONFIGS[name], in_fmt=F.IN_U8, S=1, N=24 * 4096)
    iq = F.make_input(cfg)[0]
    ch = np.ascontiguousarray(iq[:, 0])
    x = np.stack([ch, np.full_like(ch, 128)], axis=1).reshape(1, -1, 2)
    ref, _ = O.chain_run(x, F.IN_U8, cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib)
    assert np.all(ref[0].imag == 0)
    (tmp_path / "chain.bin").write_bytes(_chain_bytes(cfg))
    (tmp_path / "in.u8").write_bytes(ch.tobytes())
    n_items = len(ref[0])
    r = subprocess.run([SINK_EXE, "source_f", str(tmp_path / "chain.bin"), str(tmp_path / "in.u8"), str(tmp_path / "out.f32"), str(n_items)],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    y = np.fromfile(tmp_path / "out.f32", np.float32)
    assert y.shape == ref[0].real.shape
    nrms, peak = F.err_metrics(y, ref[0].real)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.gpu
@pytest.mark.parametrize("name,factory", [("cfg5_small", "desc"), ("cfg2_small", "desc"), ("cfg5_small", "vec"), ("cfg4_small", "vec")])
def test_gr_block_work_matches_reference_classes(gpu_lib, oracle_lib, tmp_path, name, factory):
    from oracle import oracle as O
    if not os.path.exists(EXE):
        subprocess.run(["python", os.path.join(HERE, "cpp", "build_cpp.py")], check=True)
    cfg = dict(F.CONFIGS[name], in_fmt=F.IN_U8, S=1, N=24 * 4096)
    iq = F.make_input(cfg)[0]                      # (N, 2) uint8, I then Q, as the board sends them
    x = iq.reshape(1, -1, 2)
    if name == "cfg4_small":
        ref = _reference_fed_per_block(O, oracle_lib, iq, cfg, 4096)[None]
    else:                                          # integral steps: the split into calls does not matter
        ref, _ = O.chain_run(x, F.IN_U8, cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib)
    n_items = len(ref[0])                          # whole blocks only: every input block is consumed
    (tmp_path / "chain.bin").write_bytes(_chain_bytes(cfg))
    (tmp_path / "in.u8").write_bytes(iq.tobytes())
    r = subprocess.run([EXE, str(tmp_path / "chain.bin"), str(tmp_path / "in.u8"), str(tmp_path / "out.c64"), str(n_items), factory],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "overflows 0, discarded 0" in r.stdout
    y = np.fromfile(tmp_path / "out.c64", np.complex64)
    assert y.shape == ref[0].shape
    nrms, peak = F.err_metrics(y, ref[0])
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)
