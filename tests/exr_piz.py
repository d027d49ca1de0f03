"""PIZ-compressed OpenEXR chunks, ENCODER side: test infrastructure for host/envmap_io.cpp's PIZ reader (the reference reads such maps through its
vendored tinyexr, CGL/include/CGL/tinyexr.h: DecompressPiz; its encoder half is compiled out, so the fixtures are written here and checked against
the reference's own decoder by oracle/exr_probe.cpp).  Format as OpenEXR publishes it (ImfPizCompressor / ImfWav / ImfHuf):

  chunk   = minNonZero u16, maxNonZero u16, bitmap[minNonZero .. maxNonZero], length i32, Huffman stream
  samples = per channel (header order) a plane of ny rows x nx pixels x (1 | 2) 16-bit words; the values present in the chunk are renumbered
            densely through the bitmap, every 16-bit plane is Haar-wavelet transformed in place (14-bit form when the largest number is below
            2^14, the modulo-2^16 form otherwise), and the words of all planes are Huffman coded as one sequence
  Huffman = im u32, iM u32, table bytes u32, data bits u32, 0 u32, code lengths of symbols im .. iM in 6 bits each (59..62: 2..5 zero lengths,
            63 + 8 bits: 6..261 zero lengths), padded to a byte; then the codes MSB first.  Codes are canonical: within a length in symbol order,
            the longest codes numerically first.  Symbol iM is the run marker: marker + 8-bit count repeats the previous word `count` times."""
import heapq
import struct

import numpy as np


def _wenc14(a, b):
    a = a - 65536 if a >= 32768 else a
    b = b - 65536 if b >= 32768 else b
    return ((a + b) >> 1) & 0xFFFF, (a - b) & 0xFFFF


def _wenc16(a, b):
    ao = (a + 32768) & 0xFFFF
    m = (ao + b) >> 1
    d = ao - b
    if d < 0:
        m = (m + 32768) & 0xFFFF
    return m, d & 0xFFFF


def wavelet_encode(buf, start, nx, ox, ny, oy, mx):
    """2-D Haar transform of the plane buf[start + x * ox + y * oy], in place (buf: list of ints)"""
    enc = _wenc14 if mx < (1 << 14) else _wenc16
    n = min(nx, ny)
    p, p2 = 1, 2
    while p2 <= n:
        oy1, oy2, ox1, ox2 = oy * p, oy * p2, ox * p, ox * p2
        py, ey = start, start + oy * (ny - p2)
        while py <= ey:
            px, ex = py, py + ox * (nx - p2)
            while px <= ex:
                p01, p10 = px + ox1, px + oy1
                p11 = p10 + ox1
                i00, i01 = enc(buf[px], buf[p01])
                i10, i11 = enc(buf[p10], buf[p11])
                buf[px], buf[p10] = enc(i00, i10)
                buf[p01], buf[p11] = enc(i01, i11)
                px += ox2
            if nx & p:
                p10 = px + oy1
                buf[px], buf[p10] = enc(buf[px], buf[p10])
            py += oy2
        if ny & p:
            px, ex = py, py + ox * (nx - p2)
            while px <= ex:
                p01 = px + ox1
                buf[px], buf[p01] = enc(buf[px], buf[p01])
                px += ox2
        p, p2 = p2, p2 << 1


class _Bits:
    def __init__(self):
        self.out, self.acc, self.n, self.total = bytearray(), 0, 0, 0

    def put(self, nbits, value):
        self.acc = (self.acc << nbits) | (value & ((1 << nbits) - 1)); self.n += nbits; self.total += nbits
        while self.n >= 8:
            self.n -= 8
            self.out.append((self.acc >> self.n) & 0xFF)
        self.acc &= (1 << self.n) - 1

    def flush(self):
        if self.n:
            self.out.append((self.acc << (8 - self.n)) & 0xFF)
        self.acc = self.n = 0
        return bytes(self.out)


def huffman_lengths(freq):
    """code length per symbol of a complete prefix code (plain Huffman; the format allows up to 58 bits)"""
    heap = [(f, s, None, None) for s, f in sorted(freq.items())]
    heapq.heapify(heap)
    serial = 1 << 20
    while len(heap) > 1:
        a = heapq.heappop(heap); b = heapq.heappop(heap)
        heapq.heappush(heap, (a[0] + b[0], serial, a, b)); serial += 1
    lengths = {}
    stack = [(heap[0], 0)]
    while stack:
        node, depth = stack.pop()
        if node[2] is None:
            lengths[node[1]] = max(depth, 1)
        else:
            stack.append((node[2], depth + 1)); stack.append((node[3], depth + 1))
    assert max(lengths.values()) <= 58
    return lengths


def canonical_codes(lengths):
    count = [0] * 59
    for l in lengths.values():
        count[l] += 1
    start, c = [0] * 59, 0
    for l in range(58, 0, -1):      # the longest codes take the numerically smallest values
        start[l] = c
        c = (c + count[l]) >> 1
    codes = {}
    for s in sorted(lengths):
        l = lengths[s]
        codes[s] = start[l]; start[l] += 1
    return codes


def huffman_compress(words):
    """16-bit words -> the format's Huffman stream"""
    if len(words) == 0:
        return b""
    freq = {}
    for v in words:
        freq[v] = freq.get(v, 0) + 1
    im, iM = min(freq), max(freq) + 1
    freq[iM] = 1                    # the run marker
    lengths = huffman_lengths(freq)
    codes = canonical_codes(lengths)
    tb = _Bits()
    s = im
    while s <= iM:
        if s in lengths:
            tb.put(6, lengths[s]); s += 1
            continue
        z = s
        while z <= iM and z not in lengths:
            z += 1
        run = min(z - s, 261)
        if run >= 6:
            tb.put(6, 63); tb.put(8, run - 6)
        elif run >= 2:
            tb.put(6, 59 + run - 2)
        else:
            tb.put(6, 0)
        s += run
    table = tb.flush()
    db = _Bits()
    i, n = 0, len(words)
    lm, cm = lengths[iM], codes[iM]
    while i < n:
        v = words[i]
        j = i
        while j + 1 < n and words[j + 1] == v and j - i < 255:
            j += 1
        reps = j - i                # repeats after the first
        l, c = lengths[v], codes[v]
        if l + lm + 8 < l * reps:
            db.put(l, c); db.put(lm, cm); db.put(8, reps)
        else:
            for _ in range(reps + 1):
                db.put(l, c)
        i = j + 1
    nbits = db.total
    data = db.flush()
    return struct.pack("<IIIII", im, iM, len(table), nbits, 0) + table + data


def piz_chunk(planes_per_channel):
    """planes_per_channel: per channel a uint16 array [ny, nx * size] (size 1: half, 2: the two halves of each 32-bit sample, low first).
    Returns the compressed chunk (without the y / length prefix)."""
    words = []
    geo = []
    for a in planes_per_channel:
        ny, row = a.shape
        geo.append((len(words), ny, row))
        words.extend(int(v) for v in a.reshape(-1))
    present = np.zeros(65536, bool)
    present[np.asarray(words, np.int64)] = True
    bitmap = np.packbits(present.reshape(-1, 8)[:, ::-1], axis=1).reshape(-1).copy()   # bit (v & 7) of byte (v >> 3)
    bitmap[0] &= 0xFE                                     # zero is always present and never stored
    nz = np.nonzero(bitmap)[0]
    lo, hi = (int(nz[0]), int(nz[-1])) if len(nz) else (8191, 0)
    lut = np.zeros(65536, np.int64)
    ids = np.nonzero(present | (np.arange(65536) == 0))[0]
    lut[ids] = np.arange(len(ids))
    mx = len(ids) - 1
    words = [int(lut[v]) for v in words]
    for (start, ny, row), a in zip(geo, planes_per_channel):
        size = a.size_words
        nx = row // size
        for j in range(size):
            wavelet_encode(words, start + j, nx, size, ny, row, mx)
    huf = huffman_compress(words)
    out = struct.pack("<HH", lo, hi)
    if lo <= hi:
        out += bitmap[lo:hi + 1].tobytes()
    return out + struct.pack("<i", len(huf)) + huf


class _Plane(np.ndarray):
    size_words = 1


def write_exr_piz(path, img, pixel_type, lines_per_chunk=32, force_compressed=True):
    """scan-line OpenEXR file, PIZ compression, channels B G R of type `pixel_type` (1 half, 2 float); img [h, w, 3] float32"""
    h, w, _ = img.shape

    def attr(name, typ, data):
        return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(data)) + data
    chl = b"".join(n.encode() + b"\0" + struct.pack("<iBBBBii", pixel_type, 0, 0, 0, 0, 1, 1) for n in "BGR") + b"\0"
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    hdr = (struct.pack("<II", 20000630, 2) + attr("channels", "chlist", chl) + attr("compression", "compression", bytes([4])) +
           attr("dataWindow", "box2i", box) + attr("displayWindow", "box2i", box) + attr("lineOrder", "lineOrder", b"\0") +
           attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) + attr("screenWindowCenter", "v2f", struct.pack("<ff", 0, 0)) +
           attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0")
    dt = np.float16 if pixel_type == 1 else np.float32
    size = 1 if pixel_type == 1 else 2
    chunks = []
    for y0 in range(0, h, lines_per_chunk):
        y1 = min(y0 + lines_per_chunk, h)
        planes = []
        for c in (2, 1, 0):
            a = np.ascontiguousarray(img[y0:y1, :, c], dt).view(np.uint16).reshape(y1 - y0, w * size).view(_Plane)
            a.size_words = size
            planes.append(a)
        data = piz_chunk(planes)
        raw = b"".join(np.ascontiguousarray(img[y, :, c], dt).tobytes() for y in range(y0, y1) for c in (2, 1, 0))
        if not force_compressed and len(data) >= len(raw):
            data = raw                                    # OpenEXR stores a chunk raw when the codec does not shrink it
        chunks.append(struct.pack("<ii", y0, len(data)) + data)
    off = len(hdr) + 8 * len(chunks)
    table = b""
    for r in chunks:
        table += struct.pack("<Q", off); off += len(r)
    with open(path, "wb") as f:
        f.write(hdr + table + b"".join(chunks))
