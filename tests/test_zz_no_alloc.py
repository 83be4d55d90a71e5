"""sfe_polyphase_process_noalloc (include/sfe_gpu.h section 7): the process() call that allocates nothing, against the
existing entry point sfe_polyphase_process on the same stream fed as two calls in a row.  The file name sorts last on
purpose: a failure here must not hide any other test."""
import ctypes as C

import numpy as np
import pytest

import fixtures as F

pytestmark = pytest.mark.gpu

ERR_CAP = -4
MAX_OUT = 1 << 18            # SFE_NOALLOC_MAX_OUT


def _f32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _existing(lib, klass, taps, up, blk, rate, exact, pieces):
    """sfe_{decimate,resample}_create + one sfe_polyphase_process per piece: (counts, outputs)"""
    make = lib.sfe_decimate_create if klass == F.KLASS_DECIMATE else lib.sfe_resample_create
    h = make(_f32p(taps), len(taps), up, blk)
    assert h
    lib.sfe_polyphase_set_exact(h, exact)
    outs = []
    for x in pieces:
        cap = int(len(x) / rate) + 2
        y = np.zeros(cap, np.float32)
        n = lib.sfe_polyphase_process(h, _f32p(x), len(x), _f32p(y), cap, rate)
        assert n >= 0
        outs.append(y[:n].copy())
    lib.sfe_polyphase_destroy(h)
    return outs


@pytest.mark.parametrize("klass,up,rate,exact", [(F.KLASS_RESAMPLE, 4, 1.77, 1), (F.KLASS_RESAMPLE, 4, 1.77, 0),
                                                 (F.KLASS_DECIMATE, 1, 2.5003, 1), (F.KLASS_DECIMATE, 3, 208.3333, 0)])
def test_noalloc_two_calls_match_existing_entry_point(gpu_lib, klass, up, rate, exact):
    import torch
    from simplefe_b200._lib import WalkStateC
    taps = F.taps("rs512_l3") if up == 3 else F.taps("taps31")
    H = (len(taps) - 1) // up + 1
    rng = np.random.default_rng(77)
    n1, n2 = (4096, 3001) if up != 3 else (60000, 50001)
    x1 = rng.normal(0, 0.3, n1).astype(np.float32)
    x2 = rng.normal(0, 0.3, n2).astype(np.float32)
    ref = _existing(gpu_lib, klass, taps, up, max(n1, n2), rate, exact, [x1, x2])

    torch.cuda.set_device(0)
    d_taps = torch.from_numpy(taps.copy()).cuda()
    d_x1, d_x2 = torch.from_numpy(x1).cuda(), torch.from_numpy(x2).cuda()
    cap1, cap2 = int(n1 / rate) + 2, int(n2 / rate) + 2
    d_y1 = torch.full((cap1,), 7.0, device="cuda")
    d_y2 = torch.full((cap2,), 7.0, device="cuda")

    def two_calls():
        st = WalkStateC(0, 0.0, 0)
        a = gpu_lib.sfe_polyphase_process_noalloc(klass, d_taps.data_ptr(), len(taps), up, C.byref(st), None, 0, d_x1.data_ptr(), n1,
                                                  d_y1.data_ptr(), cap1, rate, exact, None)
        # second call of the stream: its history is the tail of the first call's input buffer
        b = gpu_lib.sfe_polyphase_process_noalloc(klass, d_taps.data_ptr(), len(taps), up, C.byref(st), d_x1.data_ptr() + 4 * (n1 - H), H,
                                                  d_x2.data_ptr(), n2, d_y2.data_ptr(), cap2, rate, exact, None)
        torch.cuda.synchronize()
        return a, b

    two_calls()                                    # warm-up: module load, cudaGetSymbolAddress, the host list's reserve
    free0 = torch.cuda.mem_get_info()[0]
    a, b = two_calls()
    free1 = torch.cuda.mem_get_info()[0]
    assert free1 >= free0, "device memory was allocated by the no-alloc entry point"
    assert (a, b) == (len(ref[0]), len(ref[1])), (a, b, gpu_lib.sfe_last_error())
    for got, want in ((d_y1[:a].cpu().numpy(), ref[0]), (d_y2[:b].cpu().numpy(), ref[1])):
        if exact:
            assert np.array_equal(got, want)
        else:
            nrms, peak = F.err_metrics(got, want)
            assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)
    # nothing behind the outputs was touched
    assert float(d_y1[a:].min()) == 7.0 if a < cap1 else True
    assert float(d_y2[b:].min()) == 7.0 if b < cap2 else True


def test_noalloc_refuses_more_than_the_static_scratch(gpu_lib):
    import torch
    from simplefe_b200._lib import WalkStateC
    torch.cuda.set_device(0)
    taps = F.taps("taps31")
    d_taps = torch.from_numpy(taps.copy()).cuda()
    n_in = MAX_OUT + 4096
    d_x = torch.zeros(n_in, device="cuda")
    d_y = torch.zeros(n_in + 2, device="cuda")
    st = WalkStateC(5, 0.25, 1)
    rc = gpu_lib.sfe_polyphase_process_noalloc(F.KLASS_DECIMATE, d_taps.data_ptr(), len(taps), 1, C.byref(st), None, 0, d_x.data_ptr(), n_in,
                                               d_y.data_ptr(), n_in + 2, 1.0, 0, None)
    assert rc == ERR_CAP and "SFE_NOALLOC_MAX_OUT" in gpu_lib.sfe_last_error().decode()
    assert (st.pos, st.mu, st.held) == (5, 0.25, 1), "a refused call must leave the caller's state alone"
    # a call just under the limit still runs, and counts what the existing entry point counts
    n_ok = MAX_OUT - 4
    want = _existing(gpu_lib, F.KLASS_DECIMATE, taps, 1, n_ok, 1.0, 0, [np.zeros(n_ok, np.float32)])[0]
    st = WalkStateC(0, 0.0, 0)
    rc = gpu_lib.sfe_polyphase_process_noalloc(F.KLASS_DECIMATE, d_taps.data_ptr(), len(taps), 1, C.byref(st), None, 0, d_x.data_ptr(), n_ok,
                                               d_y.data_ptr(), n_ok + 2, 1.0, 0, None)
    torch.cuda.synchronize()
    assert rc == len(want) and rc > 0, (rc, len(want), gpu_lib.sfe_last_error())
