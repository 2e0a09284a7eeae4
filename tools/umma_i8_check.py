"""Checks the dump of tools/umma_i8_matvec.cu (tile 0: 128 permutations x 17 output words) with Python integers:
the limb columns read back from TMEM must be the exact integer row products sum_j M[i][j] * x[p][j], and the value after the
library's Montgomery reduction must be that product / 2^260 mod p (Pallas Fr), lazily reduced below 2^260.
Usage: python tools/umma_i8_check.py gpurun_out/umma_i8_dump.bin"""
import json
import struct
import sys

import numpy as np

P = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001
T, LIMBS, M = 17, 32, 128


def main(path):
    buf = open(path, "rb").read()
    o = 0
    ml = np.frombuffer(buf, np.uint8, T * T * LIMBS, o).reshape(T, T, LIMBS); o += T * T * LIMBS
    x = np.frombuffer(buf, np.uint8, M * T * LIMBS, o).reshape(M, T, LIMBS); o += M * T * LIMBS
    raw = np.frombuffer(buf, np.uint64, M * T * 10, o).reshape(M, T, 10); o += M * T * 10 * 8
    red = np.frombuffer(buf, np.uint64, M * T * 5, o).reshape(M, T, 5); o += M * T * 5 * 8
    assert o == len(buf)
    mi = [[int.from_bytes(ml[i, j].tobytes(), "little") for j in range(T)] for i in range(T)]
    bad_raw = bad_red = 0
    for p in range(M):
        xi = [int.from_bytes(x[p, j].tobytes(), "little") for j in range(T)]
        for i in range(T):
            s = sum(mi[i][j] * xi[j] for j in range(T))
            got = sum(int(raw[p, i, k]) << (52 * k) for k in range(10))
            bad_raw += got != s
            r = 0
            for k in range(5):
                d = struct.unpack("<d", struct.pack("<Q", int(red[p, i, k])))[0]
                assert d == int(d) and 0 <= d < 2 ** 53
                r += int(d) << (52 * k)
            bad_red += not (r < (1 << 260) and (r << 260) % P == s % P)
    print(json.dumps({"checked_outputs": M * T, "raw_mismatches": int(bad_raw), "reduced_mismatches": int(bad_red), "ok": bad_raw == 0 and bad_red == 0}))
    return 0 if bad_raw == 0 and bad_red == 0 else 1


if __name__ == "__main__":
    sys.exit(main(sys.argv[1]))
