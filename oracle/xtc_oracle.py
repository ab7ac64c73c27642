"""XTC (GROMACS compressed trajectory) codec, CPU restatement -- TEST INFRASTRUCTURE ONLY.

Only ``tests/`` may import this module; the product path decodes XTC on the GPU (``chacra_b200/csrc/xtc_decode.cuh``).

What it restates: the ``xdr3dfcoord`` algorithm of the xdrfile library that GROMACS, MDAnalysis (``libxdrfile2``) and
VMD's gromacsplugin (``Gromacs.h: xtc_3dfcoord``) share.  The reference reaches it through
``external/getcontacts/contact_calc/contact_utils.py:89-116`` (``molecule.read(..., 'xtc', ...)``: VMD's molfile plugin)
with the files ``chacra/scripts/process_hremd_output.py:147`` names (``state_{i}.xtc``), written by MDAnalysis in
``chacra/trajectories/process_hremd.py:300-307``.

**Parity unpinned.**  Neither the reference tree nor this image holds an XTC file, an xdrfile source or any package that
can read or write the format, so the decoder below is written from the published algorithm and checked only against
the encoder below (same author, same source of truth).  The frame layout and the bit stream:

  int32 magic = 1995 | int32 natoms | int32 step | float32 time | float32 box[9] | int32 natoms
  natoms <= 9:  float32 xyz[3 natoms]                                      (all XDR: big endian)
  otherwise:    float32 precision | int32 minint[3] | int32 maxint[3] | int32 smallidx | int32 nbytes | bytes (padded to 4)

Coordinates are integers ``round(x_nm * precision)``.  The stream (MSB first) holds per atom either three "large"
integers (offset by minint, mixed-radix packed into ``bitsize`` bits) or small differences to the previous atom
(mixed-radix in ``smallidx`` bits, radix ``magicints[smallidx]``), in runs announced by a flag bit + 5 bits; the first
two atoms of a run are stored in swapped order (water: H O H).  Output is nm; VMD scales to Angstrom (x 10).
"""
import struct

import numpy as np

MAGICINTS = [0, 0, 0, 0, 0, 0, 0, 0, 0, 8, 10, 12, 16, 20, 25, 32, 40, 50, 64, 80, 101, 128, 161, 203, 256, 322, 406, 512, 645,
             812, 1024, 1290, 1625, 2048, 2580, 3250, 4096, 5060, 6501, 8192, 10321, 13003, 16384, 20642, 26007, 32768, 41285,
             52015, 65536, 82570, 104031, 131072, 165140, 208063, 262144, 330280, 416127, 524287, 660561, 832255, 1048576,
             1321122, 1664510, 2097152, 2642245, 3329021, 4194304, 5284491, 6658042, 8388607, 10568983, 13316085, 16777216]
FIRSTIDX = 9
HEADER_BYTES = 52          # magic, natoms, step, time, box


def sizeofint(size):
    """bits needed for values 0 .. size"""
    num, bits = 1, 0
    while size >= num and bits < 32:
        bits += 1
        num <<= 1
    return bits


def sizeofints(sizes):
    """bits needed for the mixed-radix number with these radices (= bit length of their product, byte-wise)"""
    prod = 1
    for s in sizes:
        prod *= int(s)
    nbytes = max(1, (prod.bit_length() + 7) // 8)
    top = prod >> (8 * (nbytes - 1))
    num, bits = 1, 0
    while top >= num:
        bits += 1
        num *= 2
    return bits + 8 * (nbytes - 1)


class BitReader:
    """MSB-first bit stream (xdrfile receivebits)."""

    def __init__(self, data: bytes):
        self.v = int.from_bytes(data, "big")
        self.n = 8 * len(data)
        self.pos = 0

    def bits(self, k):
        if k == 0:
            return 0
        if self.pos + k > self.n:
            raise ValueError("XTC bit stream exhausted")
        out = (self.v >> (self.n - self.pos - k)) & ((1 << k) - 1)
        self.pos += k
        return out

    def ints(self, nbits, sizes):
        """three mixed-radix digits packed into nbits: whole bytes least significant first, the rest on top"""
        full, rem = divmod(nbits, 8)
        if rem == 0 and full > 0:
            full, rem = full - 1, 8
        v = 0
        for k in range(full):
            v |= self.bits(8) << (8 * k)
        if rem:
            v |= self.bits(rem) << (8 * full)
        v, n2 = divmod(v, sizes[2])
        n0, n1 = divmod(v, sizes[1])
        return [n0, n1, n2]


class BitWriter:
    def __init__(self):
        self.v, self.n = 0, 0

    def bits(self, k, value):
        assert 0 <= value < (1 << k) or k == 0
        self.v = (self.v << k) | value
        self.n += k

    def ints(self, nbits, sizes, nums):
        v = (nums[0] * sizes[1] + nums[1]) * sizes[2] + nums[2]
        assert v < (1 << nbits)
        full, rem = divmod(nbits, 8)
        if rem == 0 and full > 0:
            full, rem = full - 1, 8
        for k in range(full):
            self.bits(8, (v >> (8 * k)) & 0xff)
        if rem:
            self.bits(rem, v >> (8 * full))

    def tobytes(self):
        pad = (-self.n) % 8
        return ((self.v << pad).to_bytes((self.n + pad) // 8, "big")) if self.n else b""


def decode_ints(data: bytes, natoms, minint, maxint, smallidx):
    """the integer coordinates [natoms][3] of one frame's bit stream"""
    sizeint = [maxint[k] - minint[k] + 1 for k in range(3)]
    if (sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff:
        bitsizeint, bitsize = [sizeofint(s) for s in sizeint], 0
    else:
        bitsizeint, bitsize = None, sizeofints(sizeint)
    smaller = MAGICINTS[max(FIRSTIDX, smallidx - 1)] // 2
    smallnum = MAGICINTS[smallidx] // 2
    sizesmall = [MAGICINTS[smallidx]] * 3
    rd = BitReader(data)
    out = np.zeros((natoms, 3), np.int64)
    i, run = 0, 0
    while i < natoms:
        if bitsize == 0:
            this = [rd.bits(bitsizeint[k]) for k in range(3)]
        else:
            this = rd.ints(bitsize, sizeint)
        this = [this[k] + minint[k] for k in range(3)]
        i += 1
        flag = rd.bits(1)
        is_smaller = 0
        if flag:
            run = rd.bits(5)
            is_smaller = run % 3
            run -= is_smaller
            is_smaller -= 1
        if run > 0:
            prev = this
            for k in range(0, run, 3):
                d = rd.ints(smallidx, sizesmall)
                cur = [d[c] + prev[c] - smallnum for c in range(3)]
                i += 1
                if k == 0:          # first two atoms of a run are stored in swapped order
                    out[i - 2] = cur
                    out[i - 1] = prev
                    prev = cur
                else:
                    out[i - 1] = cur
                    prev = cur
        else:
            out[i - 1] = this
        smallidx += is_smaller
        if is_smaller < 0:
            smallnum = smaller
            smaller = MAGICINTS[smallidx - 1] // 2 if smallidx > FIRSTIDX else 0
        elif is_smaller > 0:
            smaller = smallnum
            smallnum = MAGICINTS[smallidx] // 2
        sizesmall = [MAGICINTS[smallidx]] * 3
    if i != natoms:
        raise ValueError("XTC run overshoots the atom count")
    return out


def encode_ints(ints, max_run_atoms=8, adapt=True):
    """A VALID (not size-optimal) bit stream for integer coordinates [natoms][3]: greedy runs of atoms whose differences
    fit the current small radix, the radix index stepped down / up as the differences allow (exercises every decoder
    path: large-only atoms, runs with the swapped first pair, is_smaller = -1 / 0 / +1, the unchanged-run flag 0).
    Returns (bytes, minint, maxint, smallidx0)."""
    ints = np.asarray(ints, np.int64)
    natoms = len(ints)
    minint = [int(v) for v in ints.min(0)]
    maxint = [int(v) for v in ints.max(0)]
    sizeint = [maxint[k] - minint[k] + 1 for k in range(3)]
    if (sizeint[0] | sizeint[1] | sizeint[2]) > 0xffffff:
        bitsizeint, bitsize = [sizeofint(s) for s in sizeint], 0
    else:
        bitsizeint, bitsize = None, sizeofints(sizeint)
    # starting radix: the smallest one that holds the median neighbour difference (as the real encoder's heuristic does)
    if natoms > 1:
        med = int(np.median(np.abs(np.diff(ints, axis=0)).max(1)))
    else:
        med = 0
    smallidx = FIRSTIDX
    while smallidx < len(MAGICINTS) - 1 and MAGICINTS[smallidx] // 2 <= med:
        smallidx += 1
    smallidx0 = smallidx
    smallnum = MAGICINTS[smallidx] // 2
    wr = BitWriter()
    i, prevrun = 0, -1

    def fits(delta, num):
        return all(-num <= int(d) < num for d in delta) and all(0 <= int(d) + num < 2 * num for d in delta)

    while i < natoms:
        # how many following atoms continue a run from atom i (first difference i -> i+1 is stored swapped: i - (i+1))
        k = 0
        if i + 1 < natoms and fits(ints[i] - ints[i + 1], smallnum) and MAGICINTS[smallidx] > 0:
            k = 1
            prev = ints[i]
            while k < max_run_atoms and i + 1 + k < natoms and fits(ints[i + 1 + k] - prev, smallnum):
                prev = ints[i + 1 + k]
                k += 1
        # the large coordinates: atom i+1 when a run follows (swap), else atom i
        large = ints[i + 1] if k else ints[i]
        rel = [int(large[c]) - minint[c] for c in range(3)]
        if bitsize == 0:
            for c in range(3):
                wr.bits(bitsizeint[c], rel[c])
        else:
            wr.ints(bitsize, sizeint, rel)
        run = 3 * k
        # radix adaptation for the NEXT atoms: down when this run's differences would fit the smaller radix, up when no run
        is_smaller = 0
        if adapt:
            if k == 0 and smallidx < len(MAGICINTS) - 1 and i + 1 < natoms:
                is_smaller = 1
            elif k and smallidx > FIRSTIDX:
                small2 = MAGICINTS[smallidx - 1] // 2
                seq = [ints[i] - ints[i + 1]] + [ints[i + 1 + m] - (ints[i] if m == 1 else ints[i + m]) for m in range(1, k)]
                if all(fits(d, small2) for d in seq):
                    is_smaller = -1
        if run != prevrun or is_smaller != 0:
            wr.bits(1, 1)
            wr.bits(5, run + is_smaller + 1)
            prevrun = run
        else:
            wr.bits(1, 0)
        if k:
            sizesmall = [MAGICINTS[smallidx]] * 3
            d = ints[i] - ints[i + 1]
            wr.ints(smallidx, sizesmall, [int(d[c]) + smallnum for c in range(3)])
            prev = ints[i]
            for m in range(1, k):
                d = ints[i + 1 + m] - prev
                wr.ints(smallidx, sizesmall, [int(d[c]) + smallnum for c in range(3)])
                prev = ints[i + 1 + m]
        i += 1 + k
        smallidx += is_smaller
        smallnum = MAGICINTS[smallidx] // 2
    return wr.tobytes(), minint, maxint, smallidx0


# ---- frames and files ----------------------------------------------------------------------------------------
def quantize_nm(xyz_nm, precision=1000.0):
    """xdrfile's rounding: float32 product, +-0.5, truncation"""
    lf = np.asarray(xyz_nm, np.float32) * np.float32(precision)
    return np.where(lf >= 0, (lf + np.float32(0.5)).astype(np.int64), (lf - np.float32(0.5)).astype(np.int64))


def ints_to_angstrom(ints, precision=1000.0):
    """float32(int) * (1.0f / precision) in nm, times 10.0f (VMD's molfile plugin hands Angstrom to the reference)"""
    inv = np.float32(1.0) / np.float32(precision)
    return (ints.astype(np.float32) * inv) * np.float32(10.0)


def encode_frame(xyz_angstrom, step=0, time=0.0, box_nm=None, precision=1000.0, **kw):
    xyz_nm = np.asarray(xyz_angstrom, np.float32) / np.float32(10.0)
    natoms = len(xyz_nm)
    box = np.zeros(9, np.float32) if box_nm is None else np.asarray(box_nm, np.float32).reshape(9)
    head = struct.pack(">iiif", 1995, natoms, step, time) + box.astype(">f4").tobytes() + struct.pack(">i", natoms)
    if natoms <= 9:
        return head + xyz_nm.astype(">f4").tobytes()
    data, mn, mx, si = encode_ints(quantize_nm(xyz_nm, precision), **kw)
    pad = (-len(data)) % 4
    return head + struct.pack(">f3i3iii", precision, *mn, *mx, si, len(data)) + data + b"\0" * pad


def write_xtc(path, frames_angstrom, precision=1000.0, dt=1.0, **kw):
    with open(path, "wb") as fh:
        for k, f in enumerate(frames_angstrom):
            fh.write(encode_frame(f, step=k, time=k * dt, precision=precision, **kw))


def read_xtc(path):
    """float32 [F][N][3] Angstrom, decoded on the CPU (slow: pure Python bit stream)"""
    raw = open(path, "rb").read()
    pos, frames = 0, []
    while pos < len(raw):
        magic, natoms = struct.unpack_from(">ii", raw, pos)
        if magic != 1995:
            raise ValueError("bad XTC magic %d at byte %d" % (magic, pos))
        pos += HEADER_BYTES
        (lsize,) = struct.unpack_from(">i", raw, pos)
        pos += 4
        if lsize <= 9:
            xyz = np.frombuffer(raw, ">f4", 3 * lsize, pos).astype(np.float32).reshape(lsize, 3) * np.float32(10.0)
            pos += 12 * lsize
        else:
            precision, = struct.unpack_from(">f", raw, pos)
            mn = list(struct.unpack_from(">3i", raw, pos + 4))
            mx = list(struct.unpack_from(">3i", raw, pos + 16))
            si, nbytes = struct.unpack_from(">ii", raw, pos + 28)
            pos += 36
            xyz = ints_to_angstrom(decode_ints(raw[pos:pos + nbytes], lsize, mn, mx, si), precision)
            pos += (nbytes + 3) & ~3
        frames.append(xyz)
    return np.stack(frames) if frames else np.zeros((0, 0, 3), np.float32)
