"""Build nums_b200/csrc/zig_tables.h: the 256-layer ziggurat tables of NumPy's Generator.standard_normal.

The reference draws normal blocks with numpy.random.Generator(PCG64(seed).jumped(j)).normal(...)
(nums/core/kernel/numpy_kernel.py:31-32,103-110); NumPy (a third-party dependency of the reference, 2.3.x in this
image) implements that with the Marsaglia-Tsang ziggurat in the 52-bit layout of Doornik / randmtzig:
  r = next_uint64; idx = r & 0xff; sign = (r >> 8) & 1; rabs = (r >> 9) & (2^52 - 1); x = rabs * wi[idx]
  accept x when rabs < ki[idx]; idx == 0 -> tail by two log1p draws; otherwise the wedge test against fi[].
This script recomputes the tables from the published construction (x_255 = r, equal-area layers v) and then PROBES the
installed NumPy -- by setting the PCG64 state so that the next words are chosen ones -- to pin every wi[idx] (exact,
from x = 2^51 * wi) and every ki[idx] (exact, by bisection on "one word consumed").  Values that the double-precision
recurrence does not reproduce bit-for-bit are replaced by the probed ones, so the header always states NumPy's behaviour.
fi[] only enters a comparison (never an output value); it is computed as exp(-x_i^2/2) and, when NumPy's static library
is present, cross-checked against its read-only data.
"""
import math
import os
import struct
import subprocess
import sys
import tempfile

import numpy as np
from numpy.random import PCG64, Generator

MULT = 0x2360ED051FC65DA44385DF649FCCF645
M128 = (1 << 128) - 1
MULT_INV = pow(MULT, -1, 1 << 128)
R = 3.6541528853610087963519472518
V = 0.00492867323399
TWO52 = 4503599627370496.0


def recurrence():
    ki = [0] * 256
    wi = [0.0] * 256
    fi = [0.0] * 256
    x1 = R
    wi[255] = x1 / TWO52
    fi[255] = math.exp(-0.5 * x1 * x1)
    ki[0] = int(x1 * fi[255] / V * TWO52)
    wi[0] = V / fi[255] / TWO52
    fi[0] = 1.0
    for i in range(254, 0, -1):
        x = math.sqrt(-2.0 * math.log(V / x1 + fi[i + 1]))
        ki[i + 1] = int(x / x1 * TWO52)
        wi[i] = x / TWO52
        fi[i] = math.exp(-0.5 * x * x)
        x1 = x
    ki[1] = 0
    return ki, wi, fi


class Probe:
    """A Generator whose next two 64-bit words are chosen by the caller."""

    def __init__(self):
        self.bg = PCG64(0)
        self.gen = Generator(self.bg)

    def arm(self, w1, w2):
        # post-step states with a zero high half have rot = 0 and output = low half
        inc = (w2 - w1 * MULT) & M128
        s0 = ((w1 - inc) * MULT_INV) & M128
        st = self.bg.state
        st["state"] = {"state": s0, "inc": inc}
        st["has_uint32"] = 0
        st["uinteger"] = 0
        self.bg.state = st
        self.after_one = w1

    def consumed_one(self):
        return self.bg.state["state"]["state"] == self.after_one


def word(idx, sign, rabs):
    return idx | (sign << 8) | (rabs << 9)


def probe_tables():
    p = Probe()
    # sanity: the armed words come out as chosen
    p.arm(0x0123456789ABCDEF, 0x0FEDCBA987654321)
    got = p.bg.random_raw(2)
    assert int(got[0]) == 0x0123456789ABCDEF and int(got[1]) == 0x0FEDCBA987654321, got
    wi = [0.0] * 256
    ki = [0] * 256
    for idx in range(256):
        # ki: smallest rabs that is NOT accepted on the first word (second word = all ones -> wedge/tail keeps going)
        lo, hi = 0, 1 << 52            # invariant: rabs < lo accepted at once, rabs >= hi not
        while lo < hi:
            mid = (lo + hi) // 2
            p.arm(word(idx, 0, mid), (1 << 64) - 1)
            p.gen.standard_normal()
            if p.consumed_one():
                lo = mid + 1
            else:
                hi = mid
        ki[idx] = lo
        # wi: x = rabs * wi with rabs a power of two is exact; second word 0 makes every wedge test accept
        rabs = 1 << 51
        p.arm(word(idx, 0, rabs), 0)
        x = float(p.gen.standard_normal())
        if idx == 0 and not p.consumed_one():
            raise RuntimeError("tail entered while probing wi[0]")
        wi[idx] = x / float(rabs)
    return ki, wi


def static_lib_fi():
    lib = os.path.join(os.path.dirname(np.__file__), "random", "lib", "libnpyrandom.a")
    if not os.path.exists(lib):
        return None
    with tempfile.TemporaryDirectory() as td:
        try:
            subprocess.run(["ar", "x", lib], cwd=td, check=True)
            obj = [f for f in os.listdir(td) if "distributions_distributions" in f][0]
            nm = subprocess.run(["nm", "-S", obj], cwd=td, check=True, capture_output=True, text=True).stdout
            ent = {l.split()[3]: (int(l.split()[0], 16), int(l.split()[1], 16)) for l in nm.splitlines()
                   if len(l.split()) == 4}
            sec = os.path.join(td, "ro.bin")
            subprocess.run(["objcopy", "-O", "binary", "--only-section=.rodata", obj, sec], cwd=td, check=True)
            blob = open(sec, "rb").read()
            out = {}
            for name in ("fi_double", "wi_double", "ki_double"):
                if name in ent:
                    off, size = ent[name]
                    fmt = "<256Q" if name.startswith("ki") else "<256d"
                    out[name] = list(struct.unpack(fmt, blob[off:off + size]))
            return out
        except Exception as e:   # binutils missing or layout changed: the cross-check is optional
            print("static library cross-check skipped:", e, file=sys.stderr)
            return None


def main():
    ki_r, wi_r, fi_r = recurrence()
    ki_p, wi_p = probe_tables()
    bad_k = [i for i in range(256) if ki_r[i] != ki_p[i]]
    bad_w = [i for i in range(256) if wi_r[i] != wi_p[i]]
    print(f"recurrence vs probed NumPy {np.__version__}: ki differs at {len(bad_k)} indices, wi at {len(bad_w)}",
          file=sys.stderr)
    fi = list(fi_r)
    ro = static_lib_fi()
    if ro is not None:
        if "wi_double" in ro:
            assert ro["wi_double"] == wi_p, "probed wi disagrees with the static library"
        if "ki_double" in ro:
            assert ro["ki_double"] == ki_p, "probed ki disagrees with the static library"
        if "fi_double" in ro:
            bad_f = [i for i in range(256) if ro["fi_double"][i] != fi_r[i]]
            print(f"fi: exp(-x^2/2) differs from the static library at {len(bad_f)} indices "
                  f"(max {max([abs(ro['fi_double'][i] / fi_r[i] - 1) for i in bad_f] + [0]):.2e} relative); "
                  "using the library's values", file=sys.stderr)
            fi = ro["fi_double"]
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "nums_b200", "csrc", "zig_tables.h")
    with open(out, "w") as f:
        f.write("// GENERATED by tools/make_ziggurat_tables.py -- do not edit.\n"
                "// Ziggurat tables of NumPy's Generator.standard_normal (256 layers, 52-bit rabs), pinned by probing\n"
                f"// NumPy {np.__version__}: ki/wi are exact statements of its behaviour, fi only enters a comparison.\n"
                "// Define NB_ZIG_QUAL before including (e.g. `static __device__`) to place the tables.\n"
                "#pragma once\n#include <cstdint>\n#ifndef NB_ZIG_QUAL\n#define NB_ZIG_QUAL static\n#endif\nnamespace nb {\n"
                "#define NB_ZIG_R 3.6541528853610087963519472518\n"
                "#define NB_ZIG_INV_R 0.27366123732975827203338247596\n")
        f.write("NB_ZIG_QUAL const uint64_t zig_ki[256] = {\n")
        for i in range(0, 256, 4):
            f.write("  " + ", ".join(f"0x{ki_p[j]:013X}ull" for j in range(i, i + 4)) + ",\n")
        f.write("};\nNB_ZIG_QUAL const uint64_t zig_wi_bits[256] = {   // IEEE-754 bit patterns of wi\n")
        for i in range(0, 256, 4):
            f.write("  " + ", ".join(f"0x{struct.unpack('<Q', struct.pack('<d', wi_p[j]))[0]:016X}ull"
                                     for j in range(i, i + 4)) + ",\n")
        f.write("};\nNB_ZIG_QUAL const uint64_t zig_fi_bits[256] = {   // IEEE-754 bit patterns of fi\n")
        for i in range(0, 256, 4):
            f.write("  " + ", ".join(f"0x{struct.unpack('<Q', struct.pack('<d', fi[j]))[0]:016X}ull"
                                     for j in range(i, i + 4)) + ",\n")
        f.write("};\n}  // namespace nb\n")
    print("wrote", os.path.normpath(out), file=sys.stderr)


if __name__ == "__main__":
    main()
