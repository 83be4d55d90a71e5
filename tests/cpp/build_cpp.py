"""Builds the C++ drop-in checks (test infrastructure) in-tree so that they travel to the GPU box.

tests/cpp/_bin/dropin_check        tests/cpp/dropin_check.cxx against include/libdsp/*.h + libsfegpu.so
oracle/_ref/test_blkconv_gpu       the REFERENCE's own libdsp/test/test_blkconv.cxx, compiled unchanged from
                                   /root/reference against the drop-in headers (only where the reference is mounted)
oracle/_ref/bpsk_dropin.o          examples/bpsk/bpsk.cxx compiled (not linked: libsimpleFE/libusb are out of
                                   scope) against the drop-in blkconv.h / ringbuf.h: source compatibility proof
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
INC = os.path.join(ROOT, "include", "libdsp")
LIBDIR = os.path.join(ROOT, "simplefe_b200")
BIN = os.path.join(HERE, "_bin")
REF = "/root/reference"
REFOUT = os.path.join(ROOT, "oracle", "_ref")


def _link_flags(out_dir):
    rel = os.path.relpath(LIBDIR, out_dir)
    return ["-L" + LIBDIR, "-lsfegpu", "-Wl,-rpath,$ORIGIN/" + rel]


def build(verbose=False):
    os.makedirs(BIN, exist_ok=True)
    built = []
    exe = os.path.join(BIN, "dropin_check")
    subprocess.run(["g++", "-O2", "-std=c++11", "-I" + INC, os.path.join(HERE, "dropin_check.cxx"), "-o", exe] + _link_flags(BIN),
                   check=True)
    built.append(exe)
    # GNU Radio block against stand-in gnuradio/boost headers: compile + link + run check
    exe_gr = os.path.join(BIN, "gr_block_check")
    subprocess.run(["g++", "-O2", "-std=c++11", "-pthread", "-I" + os.path.join(HERE, "stubs"), "-I" + os.path.join(ROOT, "include"),
                    os.path.join(HERE, "gr_block_check.cxx"), os.path.join(ROOT, "simplefe_b200", "gr", "gpu_source_c_impl.cc"),
                    "-o", exe_gr] + _link_flags(BIN), check=True)
    built.append(exe_gr)
    # the sink blocks and the one-channel source (same stand-in headers)
    exe_snk = os.path.join(BIN, "gr_sink_check")
    subprocess.run(["g++", "-O2", "-std=c++11", "-pthread", "-I" + os.path.join(HERE, "stubs"), "-I" + os.path.join(ROOT, "include"),
                    os.path.join(HERE, "gr_sink_check.cxx"), os.path.join(ROOT, "simplefe_b200", "gr", "gpu_sink_impl.cc"),
                    os.path.join(ROOT, "simplefe_b200", "gr", "gpu_source_f_impl.cc"), "-o", exe_snk] + _link_flags(BIN), check=True)
    built.append(exe_snk)
    if os.path.isdir(os.path.join(REF, "libdsp")):
        os.makedirs(REFOUT, exist_ok=True)
        exe2 = os.path.join(REFOUT, "test_blkconv_gpu")
        subprocess.run(["g++", "-O2", "-I" + INC, os.path.join(REF, "libdsp", "test", "test_blkconv.cxx"), "-o", exe2] + _link_flags(REFOUT),
                       check=True)
        obj = os.path.join(REFOUT, "bpsk_dropin.o")
        subprocess.run(["g++", "-O2", "-w", "-c", "-I" + INC, "-I" + os.path.join(REF, "libsimpleFE"),
                        "-I" + os.path.join(REF, "examples", "bpsk"), os.path.join(REF, "examples", "bpsk", "bpsk.cxx"), "-o", obj],
                       check=True)
        built += [exe2, obj]
    if verbose:
        print("\n".join(built))
    return built


if __name__ == "__main__":
    build(verbose=True)
