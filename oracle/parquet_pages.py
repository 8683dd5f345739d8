"""ORACLE / TEST INFRASTRUCTURE -- not part of the product path.

Pure-Python walker over an uncompressed Parquet file: Thrift compact-protocol
reader for FileMetaData / PageHeader and a scalar restatement of the
RLE/bit-packed hybrid decoder (Parquet format spec, "Encodings.md": RLE = 3,
RLE_DICTIONARY = 8, PLAIN = 0).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg may import this module.

The reference (bilalbari/spark-rapids-benchmarks) holds no decoder: its scan is
`spark.read.format(fmt).load(path)` at /root/reference/nds/nds_power.py:97-100,
executed inside Apache Spark (parquet-mr 1.10/1.12 vectorised reader, not
vendored).  This file restates the published Parquet format; the independent
cross-check is pyarrow.parquet.read_table (see tests/test_oracle_decode.py).
"""
import struct


class TReader:
    """Thrift compact protocol reader (generic: returns {field_id: value})."""

    def __init__(self, buf, pos=0):
        self.b = buf
        self.p = pos

    def byte(self):
        v = self.b[self.p]
        self.p += 1
        return v

    def varint(self):
        r = 0
        s = 0
        while True:
            c = self.byte()
            r |= (c & 0x7F) << s
            if not (c & 0x80):
                return r
            s += 7

    def zigzag(self):
        v = self.varint()
        return (v >> 1) ^ -(v & 1)

    def binary(self):
        n = self.varint()
        v = bytes(self.b[self.p:self.p + n])
        self.p += n
        return v

    def value(self, t):
        if t == 1:
            return True
        if t == 2:
            return False
        if t == 3:
            v = self.byte()
            return v - 256 if v > 127 else v
        if t in (4, 5, 6):
            return self.zigzag()
        if t == 7:
            v = struct.unpack_from('<d', self.b, self.p)[0]
            self.p += 8
            return v
        if t == 8:
            return self.binary()
        if t in (9, 10):
            h = self.byte()
            n = h >> 4
            et = h & 0xF
            if n == 15:
                n = self.varint()
            if et in (1, 2):  # list<bool>: one byte each
                return [self.byte() == 1 for _ in range(n)]
            return [self.value(et) for _ in range(n)]
        if t == 12:
            return self.struct()
        raise ValueError('thrift type %d' % t)

    def struct(self):
        out = {}
        fid = 0
        while True:
            h = self.byte()
            if h == 0:
                return out
            d = h >> 4
            t = h & 0xF
            fid = fid + d if d else self.zigzag()
            out[fid] = self.value(t)


def file_metadata(buf):
    assert buf[:4] == b'PAR1' and buf[-4:] == b'PAR1'
    flen = struct.unpack_from('<I', buf, len(buf) - 8)[0]
    return TReader(buf, len(buf) - 8 - flen).struct()


def page_header(buf, pos):
    r = TReader(buf, pos)
    h = r.struct()
    return h, r.p


def hybrid_decode(buf, pos, end, bw, n):
    """RLE/bit-packed hybrid -> list of n ints. Scalar restatement of the spec."""
    out = []
    bytew = (bw + 7) // 8
    while len(out) < n and pos < end:
        r = TReader(buf, pos)
        h = r.varint()
        pos = r.p
        if h & 1:
            groups = h >> 1
            nbytes = groups * bw
            chunk = int.from_bytes(buf[pos:pos + nbytes], 'little')
            mask = (1 << bw) - 1
            for i in range(groups * 8):
                if len(out) >= n:
                    break
                out.append((chunk >> (i * bw)) & mask)
            pos += nbytes
        else:
            cnt = h >> 1
            v = int.from_bytes(buf[pos:pos + bytew], 'little')
            pos += bytew
            out.extend([v] * min(cnt, n - len(out)))
    return out, pos


def hybrid_runs(buf, pos, end, bw, n):
    """Run structure of a hybrid stream: list of (is_packed, count)."""
    runs = []
    got = 0
    bytew = (bw + 7) // 8
    while got < n and pos < end:
        r = TReader(buf, pos)
        h = r.varint()
        pos = r.p
        if h & 1:
            c = (h >> 1) * 8
            pos += (h >> 1) * bw
        else:
            c = h >> 1
            pos += bytew
        runs.append((h & 1, c))
        got += c
    return runs


PHYS_SIZE = {1: 4, 2: 8, 4: 4, 5: 8}  # INT32, INT64, FLOAT, DOUBLE


def decode_column_chunk(buf, cmeta, max_def):
    """Decode one uncompressed column chunk to a python list (None = null).

    cmeta: thrift ColumnMetaData dict (field ids per parquet.thrift).
    Supports PLAIN and RLE_DICTIONARY data pages v1 of INT32/INT64/BYTE_ARRAY.
    """
    ptype = cmeta[1]
    nvals = cmeta[5]
    pos = cmeta[9]
    if 11 in cmeta and cmeta[11] is not None and 0 < cmeta[11] < pos:
        pos = cmeta[11]
    out = []
    dictionary = None

    def plain(pos, count):
        vals = []
        for _ in range(count):
            if ptype == 1:
                vals.append(struct.unpack_from('<i', buf, pos)[0]); pos += 4
            elif ptype == 2:
                vals.append(struct.unpack_from('<q', buf, pos)[0]); pos += 8
            elif ptype == 5:
                vals.append(struct.unpack_from('<d', buf, pos)[0]); pos += 8
            elif ptype == 6:
                ln = struct.unpack_from('<I', buf, pos)[0]; pos += 4
                vals.append(bytes(buf[pos:pos + ln])); pos += ln
            else:
                raise ValueError('physical type %d' % ptype)
        return vals, pos

    while len(out) < nvals:
        h, dpos = page_header(buf, pos)
        size = h[3]
        assert h[2] == h[3], 'compressed page'
        if h[1] == 2:  # DICTIONARY_PAGE
            dictionary, _ = plain(dpos, h[7][1])
        elif h[1] == 0:  # DATA_PAGE v1
            dh = h[5]
            n = dh[1]
            enc = dh[2]
            p = dpos
            if max_def > 0:
                ln = struct.unpack_from('<I', buf, p)[0]
                p += 4
                defs, _ = hybrid_decode(buf, p, p + ln, 1, n)
                p += ln
            else:
                defs = [1] * n
            nn = sum(defs)
            if enc in (8, 2):
                bw = buf[p]
                p += 1
                idx, _ = hybrid_decode(buf, p, dpos + size, bw, nn) if bw else ([0] * nn, p)
                vals = [dictionary[i] for i in idx]
            elif enc == 0:
                vals, _ = plain(p, nn)
            else:
                raise ValueError('encoding %d' % enc)
            it = iter(vals)
            out.extend(next(it) if d else None for d in defs)
        else:
            raise ValueError('page type %d' % h[1])
        pos = dpos + size
    return out
