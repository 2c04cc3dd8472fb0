"""Reader/writer for the .d4b named-array container (oracle/d4b_dump.h).

TEST INFRASTRUCTURE.  Fixtures under tests/golden/ are .d4b files compressed
with gzip; they were produced by oracle/make_golden.sh from the unmodified
reference (oracle/_ref/ref_driver).
"""
import gzip
import struct

import numpy as np

_DT = {0: np.float64, 1: np.int32, 2: np.int64}
_ID = {np.dtype(np.float64): 0, np.dtype(np.int32): 1, np.dtype(np.int64): 2}


def load(path):
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "rb") as f:
        buf = f.read()
    assert buf[:8] == b"D4BDUMP1", "not a d4b dump"
    pos, out = 8, {}
    while pos < len(buf):
        (nl,) = struct.unpack_from("<I", buf, pos); pos += 4
        name = buf[pos:pos + nl].decode(); pos += nl
        dt, nd = struct.unpack_from("<II", buf, pos); pos += 8
        shape = struct.unpack_from("<%dQ" % nd, buf, pos); pos += 8 * nd
        n = int(np.prod(shape)) if nd else 1
        dtype = np.dtype(_DT[dt])
        out[name] = np.frombuffer(buf, dtype=dtype, count=n, offset=pos).reshape(shape).copy()
        pos += n * dtype.itemsize
    return out


def save(path, arrays):
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "wb") as f:
        f.write(b"D4BDUMP1")
        for name, a in arrays.items():
            a = np.ascontiguousarray(a)
            nb = name.encode()
            f.write(struct.pack("<I", len(nb))); f.write(nb)
            f.write(struct.pack("<II", _ID[a.dtype], a.ndim))
            f.write(struct.pack("<%dQ" % a.ndim, *a.shape))
            f.write(a.tobytes())
