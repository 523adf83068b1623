"""The shuffle wire format of include/mrgpu.h (mr_shuffle_export), restated with numpy for the CPU tests:
per destination a segment of n records {k0, k1, count} (24 bytes) followed by the bytes of long keys.
  inline key (1..15 bytes): k0 = bytes 0..7, k1 = bytes 8..14 | len << 56 (little-endian)
  long key (16+ bytes):     k0 = offset-in-byte-block << 26 | len, k1 = 0xFF << 56 | hash32
Test infrastructure only."""
import numpy as np

from oracle import oracle as O

REC = 24


def _encode(key, long_off):
    n = len(key)
    if n <= 15:
        b = key.ljust(15, b"\0")
        return int.from_bytes(b[:8], "little"), int.from_bytes(b[8:15], "little") | (n << 56)
    return (long_off << 26) | n, (0xFF << 56) | (hash(key) & 0xFFFFFFFF)


def pack_segments(rows, world):
    """rows: oracle.Rows of one rank.  Returns (uint8 buffer, byte offsets [world+1], records [world],
    long-key bytes [world])."""
    segs, nrec, lbytes = [], [], []
    for d in range(world):
        recs, blob = [], bytearray()
        for p, k, c in zip(rows.part, rows.keys(), rows.count):
            if int(p) % world != d:
                continue
            k0, k1 = _encode(k, len(blob))
            if len(k) > 15:
                blob += k
            recs.append((k0, k1, int(c)))
        seg = np.array(recs, dtype=np.uint64).reshape(-1, 3).tobytes() + bytes(blob)
        seg += b"\0" * (-len(seg) % 16)
        segs.append(seg)
        nrec.append(len(recs))
        lbytes.append(len(blob))
    off = np.zeros(world + 1, np.int64)
    np.cumsum([len(s) for s in segs], out=off[1:])
    buf = np.frombuffer(b"".join(segs) or b"\0", dtype=np.uint8).copy()
    return buf, off, np.array(nrec, np.int64), np.array(lbytes, np.int64)


def unpack_segments(buf, off, nrec):
    out = {}
    raw = buf.tobytes()
    for s in range(len(nrec)):
        base, n = int(off[s]), int(nrec[s])
        rec = np.frombuffer(raw[base:base + n * REC], dtype=np.uint64).reshape(-1, 3)
        blob = base + n * REC
        for k0, k1, c in rec.tolist():
            if (k1 >> 56) == 0xFF:
                ln, o = k0 & ((1 << 26) - 1), k0 >> 26
                key = raw[blob + o: blob + o + ln]
            else:
                ln = k1 >> 56
                key = (k0.to_bytes(8, "little") + (k1 & ((1 << 56) - 1)).to_bytes(7, "little"))[:ln]
            out[key] = out.get(key, 0) + c
    return out
